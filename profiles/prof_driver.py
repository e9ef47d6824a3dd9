"""Short single-GPU driver for ncu captures: one warm pass + a few passes of every hot-path
kernel at the BASELINE sizes (matching b=8 512x384; GN rays N=32, 200 directed edges; tracker).
Usage: python profiles/prof_driver.py [match|gn|track|all] [reps] [gn_iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from mast3r_slam import matching  # noqa: E402
from tests import synth  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
gn_iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
H, W = 384, 512
if what in ("match", "all"):
    m = synth.make_match_inputs(8, H, W, 24, seed=1235, device=dev)
    for _ in range(reps):
        matching.match_iterative_proj(m["X11"], m["X21"], m["D11"], m["D21"])
    torch.cuda.synchronize()
if what in ("gn", "all"):
    g = synth.make_graph(N=32, seed=1236, device=dev)
    for _ in range(reps):
        T = g["Twc"].clone()
        be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                             0.003, 10.0, 0.0, 1.5, gn_iters, 0.0)
    torch.cuda.synchronize()
if what in ("track", "all"):
    d = synth.make_tracker_inputs(seed=1234, device=dev)
    for _ in range(reps):
        T = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float32, device=dev)
        be.track_ray_dist(d["Xf"], d["Xk"], d["Qk"], d["valid"], T, 0.003, 10.0, 1.345, 50, 1e-3, 1e-3)
    torch.cuda.synchronize()
print("prof_driver ok", what, be.launch_count())
