"""Design-level check of the flag protocol of the tile-dataflow Cholesky (csrc/m3r_gn_solve.cu:
chol_tile_task, chol_dataflow_kernel) — a model, not the kernel: every tile task is written down
as the sequence of flag waits, tile reads, tile writes and flag posts the kernel performs, and a
random scheduler interleaves the CTAs.  Every read must find the tile in the state the arithmetic
assumes (H = assembled normal equations, L = factored).  The GPU counterpart, which stalls the
real CTAs, is tests/test_gpu_gn.py::test_cholesky_independent_of_cta_timing.

The model also shows that the two waits added for the private TRSM are necessary: without either
one the scheduler finds an interleaving in which a CTA reads a tile in the wrong state (the first
of them is what a 15 us delay of one CTA produced on the B200: a factored tile solved twice).
"""
import random

import pytest


def tile_task(i, j, nb, own_trsm, consumed_flag):
    """Operations of chol_tile_task(i, j) in program order."""
    own = own_trsm and i == j and j > 0
    if own:  # pre-sum of the sub-diagonal tile (j, j-1), then its original values
        for k in range(j - 1):
            yield ("wait", ("tile", j, k)); yield ("wait", ("tile", j - 1, k))
            yield ("read", (j, k), "L"); yield ("read", (j - 1, k), "L")
        yield ("read", (j, j - 1), "H")
        if consumed_flag:
            yield ("post", ("consumed", j, j - 1))
    for k in range(j - 1 if own else j):
        yield ("wait", ("tile", i, k)); yield ("read", (i, k), "L")
        if i != j:
            yield ("wait", ("tile", j, k)); yield ("read", (j, k), "L")
    if own:  # private TRSM against L(j-1,j-1)
        yield ("wait", ("tile", j - 1, j - 1)); yield ("read", (j - 1, j - 1), "L")
    yield ("read", (i, j), "H")
    if i != j:
        yield ("wait", ("tile", j, j)); yield ("read", (j, j), "L")
        if own_trsm and consumed_flag and i == j + 1:
            yield ("wait", ("consumed", i, j))
    yield ("write", (i, j))
    yield ("post", ("tile", i, j))


def back_all(nb, own_trsm, wait_left_tile):
    """chol_back_all by the CTA that factored the last diagonal tile."""
    if own_trsm and wait_left_tile and nb > 1:
        yield ("wait", ("tile", nb - 1, nb - 2))
    for c in range(nb - 1, -1, -1):
        for m in range(c, nb):
            yield ("read", (m, c), "L")


def run(nb, ctas, own_trsm, consumed_flag=True, wait_left_tile=True, seed=0, back_single=True):
    """-> None, or a description of the first read that found the wrong state."""
    rng = random.Random(seed)
    tiles = [(i, j) for j in range(nb) for i in range(j, nb)]  # column-major, as the kernel walks them
    if own_trsm:
        assert ctas >= len(tiles), "the kernel enables the private TRSM only with one task per CTA"
    progs = []
    for cta in range(min(ctas, len(tiles))):
        ops = []
        for t, (i, j) in enumerate(tiles):
            if t % ctas == cta:
                ops += list(tile_task(i, j, nb, own_trsm, consumed_flag))
        if back_single and (len(tiles) - 1) % ctas == cta:
            ops += list(back_all(nb, own_trsm, wait_left_tile))
        progs.append(ops)
    state = {t: "H" for t in tiles}
    flags = set()
    pc = [0] * len(progs)
    weight = [rng.choice((1, 1, 1, 30, 1000)) for _ in progs]  # some CTAs are starved for long stretches
    while True:
        ready = [c for c in range(len(progs)) if pc[c] < len(progs[c]) and
                 not (progs[c][pc[c]][0] == "wait" and progs[c][pc[c]][1] not in flags)]
        if not ready:
            assert all(pc[c] == len(progs[c]) for c in range(len(progs))), "deadlock"
            return None
        c = rng.choices(ready, [1.0 / weight[k] for k in ready])[0]
        op = progs[c][pc[c]]
        pc[c] += 1
        if op[0] == "read" and state[op[1]] != op[2]:
            return f"CTA {c} read tile {op[1]} in state {state[op[1]]}, expected {op[2]}"
        if op[0] == "write":
            state[op[1]] = "L"
        if op[0] == "post":
            flags.add(op[1])


@pytest.mark.parametrize("nb", [1, 2, 3, 7, 8])
def test_private_trsm_protocol_has_no_bad_interleaving(nb):
    ctas = nb * (nb + 1) // 2
    for seed in range(300):
        assert run(nb, ctas, own_trsm=True, seed=seed) is None


@pytest.mark.parametrize("nb,ctas", [(2, 1), (5, 4), (9, 45), (12, 7), (20, 16)])
def test_plain_protocol_any_grid(nb, ctas):
    """No private TRSM (nb > 8 or fewer CTAs than tiles): several tasks per CTA, walked in column-major order."""
    for seed in range(100):
        assert run(nb, ctas, own_trsm=False, seed=seed, back_single=nb <= 16) is None


def test_both_added_waits_are_necessary():
    nb, ctas = 4, 10
    found = [run(nb, ctas, own_trsm=True, consumed_flag=False, seed=s) for s in range(300)]
    assert any(f and "expected H" in f for f in found), "model never delays a diagonal owner enough"
    found = [run(nb, ctas, own_trsm=True, wait_left_tile=False, seed=s) for s in range(300)]
    assert any(f and f.endswith("state H, expected L") and "(3, 2)" in f for f in found)
