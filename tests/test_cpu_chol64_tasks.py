"""Design-level checks of the macro-tile Cholesky's task list (csrc/m3r_gn_chol64.cu), on the
list the library itself builds (m3r_debug_chol64_tasklist runs the device kernel's enumeration on
the host).  No GPU.

  * coverage: every product L(I,k) L(J,k)^T of the left-looking factorisation is in exactly one
    task (the chain CTA's private last term of a diagonal tile excepted), chunk indices count up;
  * order: every task's inputs are produced by earlier list entries or by the chain, and the chain
    steps those inputs need depend only on earlier entries — the argument that makes blocking
    waits inside tasks deadlock-free;
  * a scheduler model: CTAs that take the next entry with one counter and block on their inputs
    always finish, for any CTA count >= 1 and any timing."""
import ctypes as C
import random

import numpy as np
import pytest

KC = 4


@pytest.fixture(scope="module")
def tasklist(backends):
    def get(nbm):
        n = backends._lib.m3r_debug_chol64_tasklist(nbm, None, 0)
        assert n >= 0, "enumeration and workspace formula disagree"
        buf = np.zeros((max(n, 1), 4), dtype=np.int32)
        assert backends._lib.m3r_debug_chol64_tasklist(nbm, buf.ctypes.data_as(C.c_void_p), n) == n
        out = []
        for a, b, kind, kc in buf[:n].tolist():
            out.append(dict(I=a & 0xffff, J=a >> 16, k0=b & 0xffff, k1=b >> 16, kind=kind, kc=kc))
        return out
    return get


TAIL_CHAIN = 2


RECENT, KC_OLD = 4, 16


def _nch(R, t):  # chunks of [0, R) from the end: a tail of t, RECENT / KC chunks of KC, then chunks of KC_OLD
    if R <= 0:
        return 0
    n, e, d = 1, R - t, 0
    while e > 0:
        st = KC if d < RECENT else KC_OLD
        e -= st; d += st; n += 1
    return n


def _nch_diag(J):  # list tasks cover [0, J-2); the chain adds k = J-2 and k = J-1 itself
    return _nch(J - 2, TAIL_CHAIN)


def _nch_sub(J):  # list tasks cover [0, J-1); the chain adds the last term (k = J-1) itself
    return _nch(J - 1, TAIL_CHAIN)


@pytest.mark.parametrize("nbm", [1, 2, 3, 5, 9, 13, 20, 56])
def test_coverage_and_chunk_order(tasklist, nbm):
    tasks = tasklist(nbm)
    by_tile = {}
    for pos, t in enumerate(tasks):
        assert 0 <= t["J"] <= t["I"] < nbm and 0 <= t["k0"] <= t["k1"] <= t["J"]
        by_tile.setdefault((t["I"], t["J"]), []).append((pos, t))
    for J in range(nbm):
        for I in range(J, nbm):
            ts = by_tile.get((I, J), [])
            # chunk indices 0, 1, 2, ... in list order, ranges contiguous from 0
            assert [t["kc"] for _, t in ts] == list(range(len(ts)))
            k = 0
            for _, t in ts:
                assert t["k0"] == k
                k = t["k1"]
            if I == J:       # the chain adds k = J-2 and k = J-1 itself
                assert k == max(J - 2, 0) and all(t["kind"] == 0 for _, t in ts) and len(ts) == _nch_diag(J)
            elif I == J + 1:  # [0, J-1) by update tasks; the chain adds k = J-1 and does the TRSM
                assert k == max(J - 1, 0) and all(t["kind"] == 0 for _, t in ts) and len(ts) == _nch_sub(J)
            else:             # exactly one final task, last, covering the tail of [0, J)
                assert k == J and [t["kind"] for _, t in ts] == [0] * (len(ts) - 1) + [1]
            assert all(t["k1"] - t["k0"] <= KC_OLD for _, t in ts)
            assert all(t["k1"] - t["k0"] <= KC for _, t in ts if t["k1"] > J - 1 - RECENT)
            if I <= J + 1 and ts:  # chain tiles: the chunk released last is short
                assert ts[-1][1]["k1"] - ts[-1][1]["k0"] <= TAIL_CHAIN


def _chain_needs(J):
    """(tile, chunks) whose progress the chain waits for before step J."""
    need = []
    if _nch_diag(J):
        need.append(((J, J), _nch_diag(J)))
    if J >= 1 and _nch_sub(J - 1):
        need.append(((J, J - 1), _nch_sub(J - 1)))
    return need


def _chain_needs_final(J):
    """Published tile the chain needs before step J: L(J,J-2), for the last term of tile (J,J-1)."""
    return (J, J - 2) if J >= 2 else None


@pytest.mark.parametrize("nbm", [3, 9, 20, 56])
def test_list_is_a_topological_order(tasklist, nbm):
    tasks = tasklist(nbm)
    pos_chunk = {}   # (I, J, kc) -> list position
    pos_final = {}   # (I, J) -> list position of the task that publishes L(I,J)
    for p, t in enumerate(tasks):
        pos_chunk[(t["I"], t["J"], t["kc"])] = p
        if t["kind"] == 1:
            pos_final[(t["I"], t["J"])] = p
    # last list position the chain needs before it can finish step J (steps run in order)
    chain_after = []
    worst = -1
    for J in range(nbm):
        for (tile, n) in _chain_needs(J):
            worst = max(worst, pos_chunk[(tile[0], tile[1], n - 1)])
        if _chain_needs_final(J):
            worst = max(worst, pos_final[_chain_needs_final(J)])
        chain_after.append(worst)

    def producer_pos(I, k):  # list position after which L(I,k) exists
        if I == k or I == k + 1:
            return chain_after[I]  # chain step I publishes L(I,I) and L(I,I-1)
        return pos_final[(I, k)]

    for p, t in enumerate(tasks):
        if t["kc"] > 0:
            assert pos_chunk[(t["I"], t["J"], t["kc"] - 1)] < p
        for k in range(t["k0"], t["k1"]):
            assert producer_pos(t["I"], k) < p and producer_pos(t["J"], k) < p, (p, t, k)
        if t["kind"] == 1:
            assert chain_after[t["J"]] < p  # the TRSM waits for L(J,J)


@pytest.mark.parametrize("nbm,ctas,seed", [(9, 1, 0), (9, 3, 1), (20, 7, 2), (20, 40, 3), (56, 147, 4)])
def test_scheduler_model_always_finishes(tasklist, nbm, ctas, seed):
    """Workers take the next list entry and may only finish it when its inputs exist; the chain is a
    worker of its own.  Random interleaving; no step may leave everybody blocked."""
    rng = random.Random(seed)
    tasks = tasklist(nbm)
    ready = set()        # published tiles (I, J)
    progress = {}        # tile -> chunks applied
    nxt = 0
    holding = [None] * ctas
    chain_J = 0
    done_tasks = 0

    def inputs_ready(t):
        return all((t["I"], k) in ready and (t["J"], k) in ready for k in range(t["k0"], t["k1"])) and \
            progress.get((t["I"], t["J"]), 0) == t["kc"] and (t["kind"] == 0 or (t["J"], t["J"]) in ready)

    while chain_J < nbm or done_tasks < len(tasks):
        moved = False
        order = list(range(ctas + 1))
        rng.shuffle(order)
        for w in order:
            if w == ctas:  # the chain
                fin = _chain_needs_final(chain_J)
                if chain_J < nbm and all(progress.get(tile, 0) == n for tile, n in _chain_needs(chain_J)) and \
                        (fin is None or fin in ready):
                    if chain_J >= 1:
                        ready.add((chain_J, chain_J - 1))
                    ready.add((chain_J, chain_J))
                    chain_J += 1
                    moved = True
                continue
            if holding[w] is None and nxt < len(tasks):
                holding[w] = tasks[nxt]
                nxt += 1
                moved = True
            t = holding[w]
            if t is not None and inputs_ready(t) and rng.random() < 0.7:
                if t["kind"] == 0:
                    progress[(t["I"], t["J"])] = t["kc"] + 1
                else:
                    ready.add((t["I"], t["J"]))
                holding[w] = None
                done_tasks += 1
                moved = True
        if not moved:
            blocked = [t for t in holding if t is not None and not inputs_ready(t)]
            fin = _chain_needs_final(chain_J) if chain_J < nbm else None
            assert len(blocked) < sum(t is not None for t in holding) or chain_J < nbm and all(
                progress.get(tile, 0) == n for tile, n in _chain_needs(chain_J)) and (fin is None or fin in ready), \
                f"deadlock at chain step {chain_J}, {done_tasks}/{len(tasks)} tasks done"
    assert len(ready) == nbm * (nbm + 1) // 2
