"""Debug aid: one GN iteration on graphs that take the macro-tile Cholesky, solver status and the
error against a numpy fp64 solve of the same per-edge blocks.  Usage (GPU box): python profiles/debug_chol64.py"""
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
cases = [("sparse150", 150, {1: None, 2: 148, "random": 2}, 48, 64, 1400),
         ("dense150", 150, {1: None, 2: 148, 7: 75, "random": 50}, 12, 16, 1390),
         ("sparse100", 100, {1: None, 2: 98, "random": 2}, 48, 64, 1350),
         ("sparse260", 260, {1: None, 2: 258, "random": 2}, 24, 32, 1351)]
for name, N, counts, H, W, seed in cases:
    g = synth.make_graph(N=N, counts=counts, H=H, W=W, seed=seed, device=dev)
    prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
    for rep in range(2):
        Twc = g["Twc"].clone()
        s = be.GNSolver(prm, Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
        blocks = s.accumulate().cpu().numpy().copy()
        dx = s.solve_step(0.0).cpu().numpy().copy()
        torch.cuda.synchronize()
        st = torch.empty(4, dtype=torch.int32, device=dev)
        import ctypes as C
        be._lib.m3r_gn_status.restype = C.c_void_p
        ptr = be._lib.m3r_gn_status(C.c_void_p(s.ws.data_ptr()))
        off = ptr - s.ws.data_ptr()
        status = s.ws[off:off + 16].view(torch.int32).cpu().tolist()
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
        ev = np.linalg.eigvalsh(Hd)
        bad = np.abs(dx - ref).max(axis=1)
        print(f"{name} rep {rep}: status {status} |dx-ref| max {np.abs(dx - ref).max():.2e} |ref| max {np.abs(ref).max():.2e} "
              f"eig [{ev[0]:.2e}, {ev[-1]:.2e}] worst rows {np.argsort(bad)[-3:].tolist()}")
