"""GN rays accumulate+solve on the config-3 graph with (a) the generator's matches (96 % usable) and
(b) spatially coherent dropouts (about half of every edge's matches invalid in blobs ~30 px wide),
the situation of real maps (occlusions, sky, low-confidence regions).  Usage: python profiles/time_gn_sparse.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda", 0)
g = synth.make_graph(N=32, seed=1236, device=dev)
M, n = g["idx"].shape
gen = torch.Generator(device=dev).manual_seed(5)
blob = torch.cat([synth.smooth_noise((50, 1, synth.H0, synth.W0), 12.0, gen, dev) for _ in range(M // 50)]).view(M, n, 1)
valid_sparse = (g["valid"] & (blob > 0.0)).contiguous()


def run(valid):
    def once():
        T = g["Twc"].clone()
        be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], valid, g["Q"], 0.003, 10.0, 0.0, 1.5, 10, 0.0)
    for _ in range(3):
        once()
    a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(10):
        once()
    z.record(); torch.cuda.synchronize()
    return a.elapsed_time(z) / 10


print(f"usable matches: dense {float(g['valid'].float().mean()):.3f}, sparse {float(valid_sparse.float().mean()):.3f}")
print(f"10 GN iterations: dense {run(g['valid']):.3f} ms, coherent dropouts {run(valid_sparse):.3f} ms")
