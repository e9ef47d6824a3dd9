"""Timeline of the tile-dataflow Cholesky (m3r_debug_chol_trace) for a GN-rays graph.
Usage (GPU box): python profiles/chol_trace.py [N] [out.txt]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda", 0)
H, W = (384, 512) if N <= 64 else (48, 64)
counts = None if N == 32 else {1: None, 2: N - 2, 5: N - 5, "random": N // 4}
g = synth.make_graph(N=N, counts=counts, H=H, W=W, seed=1236, device=dev)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
s = be.GNSolver(prm, g["Twc"].clone(), g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
for _ in range(3):
    s.accumulate(); s.solve_step(0.0)
torch.cuda.synchronize()
nb = (7 * (N - 1) + 1 + 31) // 32
big = nb > 16  # macro-tile kernel (m3r_gn_chol64.cu): slots = chain steps, then list tasks, then back tasks
if big:
    nbm = (7 * (N - 1) + 1 + 63) // 64
    nb = 2 * nbm
    cap = nbm + be._lib.m3r_debug_chol64_tasklist(nbm, None, 0) + nb + 4 * nbm + 8 * nbm + 8
else:
    cap = nb * (nb + 1) // 2 + nb
buf = torch.zeros((cap, 4), dtype=torch.int64, device=dev)
be._lib.m3r_debug_chol_trace(buf.data_ptr(), cap)
s.accumulate()
torch.cuda.synchronize()
a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
a.record(); s.solve_step(0.0); z.record()
torch.cuda.synchronize()
be._lib.m3r_debug_chol_trace(None, 0)
t = buf.cpu().numpy()
if len(sys.argv) > 2:
    import numpy as _np
    _np.save(sys.argv[2] + ".npy", t)
t0 = t[t[:, 1] > 0, 1].min()
lines = [f"N={N} nb={nb} solve_step {a.elapsed_time(z) * 1e3:.1f} us (event); kernel span "
         f"{(t[:, 3].max() - t0) / 1e3:.1f} us"]
kinds = {0: "potrf", 1: "trsm ", 2: "back ", 3: "chunk", 4: "team1a", 5: "team1b", 6: "team0a", 7: "team0b"}
if big:  # summary by kind: busy time between inputs-ready and end, and waiting before it
    tt = t[t[:, 3] > 0]
    for k, name in kinds.items():
        sel = tt[(tt[:, 0] >> 32) == k]
        if len(sel):
            lines.append(f"{name}: {len(sel)} tasks, work {((sel[:, 3] - sel[:, 2]).sum()) / 1e3:.0f} us total "
                         f"(mean {((sel[:, 3] - sel[:, 2]).mean()) / 1e3:.2f}), wait for inputs "
                         f"{((sel[:, 2] - sel[:, 1]).sum()) / 1e3:.0f} us total (mean {((sel[:, 2] - sel[:, 1]).mean()) / 1e3:.2f})")
    pot = tt[(tt[:, 0] >> 32) == 0]
    lines.append(f"factorisation span {(pot[:, 3].max() - t0) / 1e3:.1f} us; back substitution "
                 f"{(tt[:, 3].max() - pot[:, 3].max()) / 1e3:.1f} us")
order = sorted(range(cap), key=lambda r: t[r, 3])
for r in order:
    if t[r, 3] == 0:
        continue
    k, i, j = int(t[r, 0] >> 32), int((t[r, 0] >> 16) & 0xFFFF), int(t[r, 0] & 0xFFFF)
    if k >= 8:
        continue
    if big and (k == 3 or (k == 1 and i > j + 2) or (k == 2 and i % 8) or (k >= 4 and not 18 <= i <= 24)):
        continue
    if k >= 4:
        lines.append(f"{kinds[k]} ({i:3d}) " + "  ".join(f"{(t[r, q] - t0) / 1e3:8.2f}" for q in (1, 2, 3)))
        continue
    if (not big) and nb > 16 and not (k != 1 or i == j + 1):
        continue
    lines.append(f"{kinds[k]} ({i:3d},{j:3d}) begin {(t[r, 1] - t0) / 1e3:8.2f}  ready {(t[r, 2] - t0) / 1e3:8.2f}  "
                 f"end {(t[r, 3] - t0) / 1e3:8.2f}  work {(t[r, 3] - t[r, 2]) / 1e3:6.2f} us")
out = "\n".join(lines)
print(out)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(out + "\n")
