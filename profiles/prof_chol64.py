"""ncu driver for the macro-tile Cholesky: a GN-rays graph of N keyframes (tiny images: the solve does
not depend on the image size), two accumulate + solve_step iterations.
Usage (GPU box): python profiles/prof_chol64.py [N]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda", 0)
g = synth.make_graph(N=N, counts={1: None, 2: N - 2, 5: N - 5, "random": N // 4}, H=48, W=64, seed=1236, device=dev)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
s = be.GNSolver(prm, g["Twc"].clone(), g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
for _ in range(2):
    s.accumulate(); s.solve_step(0.0)
torch.cuda.synchronize()
print("prof_chol64 ok", N, be.launch_count())
