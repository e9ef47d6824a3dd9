"""Robustness sweep of the Cholesky kernels over system sizes (every panel-count parity, padded and exact
dimensions, both kernels): one GN iteration against a numpy fp64 solve of the same per-edge blocks, twice for
bit-reproducibility.  Usage (GPU box): python profiles/fuzz_chol64.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda", 0)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
worst = 0.0
sizes = [2, 3, 5, 9, 10, 14, 19, 23, 24, 25, 27, 28, 32, 33, 37, 41, 42, 46, 51, 55, 56, 64, 65, 73, 74, 83, 92, 101, 110,
         119, 128, 129, 137, 147, 165, 183, 184, 201, 220, 239, 256, 257, 275, 300, 330]
for N in sizes:
    rng = np.random.RandomState(N)
    counts = {1: None}
    if N > 3:
        counts[2] = int(rng.randint(1, N - 2))
    if N > 9:
        counts[int(rng.randint(3, 8))] = int(rng.randint(1, N - 8))
        counts["random"] = int(rng.randint(1, max(2, N // 3)))
    g = synth.make_graph(N=N, counts=counts, H=12, W=16, seed=7000 + N, device=dev)
    out = []
    for rep in range(2):
        s = be.GNSolver(prm, g["Twc"].clone(), g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
        blocks = s.accumulate().cpu().numpy().copy()
        out.append(s.solve_step(0.0).clone())
    torch.cuda.synchronize()
    assert torch.equal(out[0], out[1]), f"N={N}: not reproducible"
    dx = out[0].cpu().numpy()
    D = 7 * (N - 1)
    Hd, bd = np.zeros((D, D)), np.zeros(D)
    ii, jj = g["ii"].cpu().numpy(), g["jj"].cpu().numpy()
    for e in range(g["M"]):
        S, v = blocks[e, :49].reshape(7, 7), blocks[e, 49:56]
        i, j = int(ii[e]) - 1, int(jj[e]) - 1
        si, sj = slice(7 * i, 7 * i + 7), slice(7 * j, 7 * j + 7)
        if i >= 0:
            Hd[si, si] += S; bd[si] -= v
        if j >= 0:
            Hd[sj, sj] += S; bd[sj] += v
        if i >= 0 and j >= 0:
            Hd[si, sj] -= S; Hd[sj, si] -= S
    ref = (-np.linalg.solve(Hd, bd)).reshape(N - 1, 7)
    err = np.abs(dx - ref).max() / max(1.0, np.abs(ref).max())
    worst = max(worst, err)
    panels = (D + 1 + 31) // 32
    print(f"N={N:4d} panels32={panels:3d} {'macro' if panels > 5 else '32x32'}: rel err {err:.1e}  |dx| max {np.abs(ref).max():.2e}", flush=True)
    assert err < 2e-6, f"N={N}: {err}"
print(f"fuzz ok: {len(sizes)} sizes, worst relative error {worst:.1e}")
