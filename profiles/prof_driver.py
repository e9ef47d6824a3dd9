"""Short single-GPU driver for ncu captures: one warm pass + a few passes of every hot-path
kernel at the BASELINE sizes (matching b=8 512x384; GN rays N=32, 200 directed edges; tracker).
`all` also runs GN calib (N=64), the calibrated tracker and the two front-end kernels once each.
`render` (also part of `all`) draws 24 keyframes with both viewer programs into a 1024x768 buffer.
Usage: python profiles/prof_driver.py [match|gn|track|render|all] [reps] [gn_iters]"""
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
if what == "all":
    g = synth.make_graph(N=64, counts={1: None, 2: 62, 5: 59, "random": 16}, seed=1237, device=dev)
    Xs = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])
    T = g["Twc"].clone()
    be.gauss_newton_calib(T, Xs, g["Cs"], g["K"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                          H, W, -10, 1e-6, 1.0, 10.0, 0.0, 1.5, 1, 0.0)
    d = synth.make_tracker_calib_inputs(H, W, seed=1234, device=dev)
    from mast3r_slam.tracker import FrameTracker
    from mast3r_slam.sim3 import Sim3
    trk = FrameTracker(None, None, dev)
    I = Sim3(torch.tensor([[0, 0, 0, 0, 0, 0, 1, 1.0]], device=dev))
    uv = torch.stack(torch.meshgrid(torch.arange(W, device=dev), torch.arange(H, device=dev), indexing="xy"), -1)
    meas = torch.cat((uv.reshape(-1, 2).float(), torch.log(d["Xk"][..., 2:3])), -1)
    try:
        trk.opt_pose_calib_sim3(d["Xf"], d["Xk"], I, Sim3(I.data.clone()), d["Qk"], d["valid"], meas,
                                d["Xk"][..., 2:3] > 1e-6, d["K"], (H, W))
    except RuntimeError as e:
        print("calib tracker:", e)
    E, n = 8, H * W
    idx = torch.randint(0, n, (E, n), device=dev)
    vm = torch.rand(E, n, 1, device=dev) > 0.1
    Qa, Qb = torch.rand(E, n, 1, device=dev) + 1, torch.rand(E, n, 1, device=dev) + 1
    be.edge_build(idx, vm, Qa, Qb, 1.5)
    X0, C0 = torch.randn(n, 3, device=dev), torch.rand(n, 1, device=dev) + 0.5
    be.pointmap_update(X0, C0, torch.randn(n, 3, device=dev), torch.rand(n, 1, device=dev) + 0.5,
                       "weighted_pointmap", torch.tensor([0.1, 0, 0, 0, 0, 0, 1, 1.0], device=dev))
    torch.cuda.synchronize()
if what in ("render", "all"):
    sc = synth.make_render_scene(K=24, seed=1242, device=dev)
    fb = torch.empty((768, 1024), dtype=torch.int64, device=dev)
    for _ in range(reps):
        for program, kw in (("surfelmap", dict(radius=0.004)), ("trianglemap", dict())):
            be.render_clear(fb)
            be.render_pointmaps(program, sc["X"], sc["C"], sc["img"], sc["models"], sc["camera"], sc["proj"], fb,
                                be.render_opts(**kw))
    be.render_resolve(fb)
    torch.cuda.synchronize()
print("prof_driver ok", what, be.launch_count())
