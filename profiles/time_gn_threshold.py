"""Solve-step time of the two Cholesky kernels at several system sizes (which one should take which size).
Run once per value of M3R_GN_SMALL_MAX_NB (read at library load): e.g.
  M3R_GN_SMALL_MAX_NB=1 python profiles/time_gn_threshold.py   # macro-tile kernel everywhere
  M3R_GN_SMALL_MAX_NB=999 python profiles/time_gn_threshold.py # 32x32 dataflow kernel everywhere"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda", 0)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
out = []
for N in [int(v) for v in os.environ.get("M3R_SIZES", "16,32,48,64,75,100,150,260").split(",")]:
    g = synth.make_graph(N=N, counts={1: None, 2: N - 2, 5: N - 5, "random": N // 4}, H=48, W=64, seed=1236, device=dev)
    s = be.GNSolver(prm, g["Twc"].clone(), g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
    for _ in range(3):
        s.accumulate(); s.solve_step(0.0)
    torch.cuda.synchronize()
    ts = []
    for _ in range(20):
        s.accumulate()
        a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
        a.record(); s.solve_step(0.0); z.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(z) * 1e3)
    ts.sort()
    out.append(f"N={N:4d} nb32={(7 * (N - 1) + 32) // 32:3d}: solve_step median {ts[len(ts) // 2]:7.1f} us  min {ts[0]:7.1f} us")
print(f"M3R_GN_SMALL_MAX_NB={os.environ.get('M3R_GN_SMALL_MAX_NB', '(default)')}")
print("\n".join(out))
