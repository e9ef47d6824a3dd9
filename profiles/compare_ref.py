"""Same-box comparison for BASELINE.md section 4: the UNMODIFIED reference extension
(oracle/_ref/mast3r_slam_backends_ref.so, built for sm_100 by oracle/build_ref.py) against
libm3r_b200.so on identical synthetic inputs, BASELINE configs 2-4.  Measurement tooling (it
executes oracle/_ref as the incumbent being timed, never as part of the product).

Timing protocol: CUDA events on the call's stream, warm-up + timed repetitions,
median and min; GN poses cloned per repetition outside the timed region; delta_thresh = 0.
Usage (GPU box):  python profiles/compare_ref.py [out.json]
"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from mast3r_slam import matching  # noqa: E402
from oracle import build_ref  # noqa: E402
from tests import synth  # noqa: E402

H, W, FD = 384, 512, 24
N_PIX = H * W
dev = torch.device("cuda", 0)
ref = build_ref.load_ref()
if ref is None:
    raise SystemExit("oracle/_ref/mast3r_slam_backends_ref.so not present")


def timed(fn, setup=None, warm=5, reps=20):
    ts = []
    for i in range(warm + reps):
        args = setup() if setup else ()
        a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn(*args)
        z.record()
        torch.cuda.synchronize()
        if i >= warm:
            ts.append(a.elapsed_time(z))
    return {"median_ms": statistics.median(ts), "min_ms": min(ts)}


# --- the reference's own Python glue around its two kernels (matching.py:25-90, image.py:5-38),
# restated here over the reference extension so the incumbent is timed at the same interface.
def ref_img_gradient(img):
    b, c, h, w = img.shape
    k = torch.tensor([[-3.0, 0, 3], [-10, 0, 10], [-3, 0, 3]], device=img.device) / 32.0
    kx = k.view(1, 1, 3, 3).repeat(c, 1, 1, 1)
    ky = k.t().contiguous().view(1, 1, 3, 3).repeat(c, 1, 1, 1)
    p = F.pad(img, (1, 1, 1, 1), mode="reflect")
    return F.conv2d(p, kx, groups=c), F.conv2d(p, ky, groups=c)


def ref_match_iterative_proj(X11, X21, D11, D21):
    b, h, w, _ = X11.shape
    rays = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = ref_img_gradient(rays)
    img = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts = F.normalize(X21.view(b, -1, 3), dim=-1)
    idx0 = torch.arange(h * w, device=X11.device)[None].repeat(b, 1)
    p_init = torch.stack((idx0 % w, idx0 // w), -1).float()
    p1, conv = ref.iter_proj(img, pts, p_init, 10, 1e-8, 1e-6)
    p1 = p1.long()
    bi = torch.arange(b, device=X11.device)[:, None]
    dist = torch.linalg.norm(X11[bi, p1[..., 1], p1[..., 0]] - X21.view(b, -1, 3), dim=-1)
    valid = conv & (dist < 0.1)
    (p1,) = ref.refine_matches(D11.half(), D21.view(b, h * w, -1).half(), p1, 3, 5)
    return p1[..., 0] + w * p1[..., 1], valid.unsqueeze(-1)


out = {"gpu": torch.cuda.get_device_name(0)}

# ---------------------------------------------------------------- config 2: matching
for b in (1, 2, 8):
    m = synth.make_match_inputs(b, H, W, FD, seed=1235, device=dev)
    X11, X21, D11, D21 = m["X11"], m["X21"], m["D11"], m["D21"]
    rays, D11h = be.match_prep(X11, D11)
    pts = F.normalize(X21.view(b, -1, 3), dim=-1).contiguous()
    idx0 = torch.arange(N_PIX, device=dev)[None].repeat(b, 1)
    p_init = torch.stack((idx0 % W, idx0 // W), -1).float().contiguous()
    D21h = D21.view(b, N_PIX, FD).half().contiguous()
    p_ref, _ = ref.iter_proj(rays, pts, p_init, 10, 1e-8, 1e-6)
    p1 = p_ref.long().contiguous()
    r = {}
    r["iter_proj_ref"] = timed(lambda: ref.iter_proj(rays, pts, p_init, 10, 1e-8, 1e-6))
    r["iter_proj_b200"] = timed(lambda: be.iter_proj(rays, pts, p_init, 10, 1e-8, 1e-6))
    r["refine_ref"] = timed(lambda: ref.refine_matches(D11h, D21h, p1, 3, 5))
    r["refine_b200"] = timed(lambda: be.refine_matches(D11h, D21h, p1, 3, 5))
    r["call_ref"] = timed(lambda: ref_match_iterative_proj(X11, X21, D11, D21))
    r["call_b200"] = timed(lambda: matching.match_iterative_proj(X11, X21, D11, D21))
    (a,) = ref.refine_matches(D11h, D21h, p1, 3, 5)
    (c,) = be.refine_matches(D11h, D21h, p1, 3, 5)
    r["refine_bit_exact"] = bool(torch.equal(a, c))
    i_ref, v_ref = ref_match_iterative_proj(X11, X21, D11, D21)
    i_new, v_new = matching.match_iterative_proj(X11, X21, D11, D21)
    r["call_idx_agree"] = float((i_ref == i_new).float().mean())
    r["call_valid_agree"] = float((v_ref == v_new).float().mean())
    for k in ("iter_proj", "refine", "call"):
        r[f"{k}_gpix_ref"] = b * N_PIX / (r[f"{k}_ref"]["median_ms"] * 1e-3) / 1e9
        r[f"{k}_gpix_b200"] = b * N_PIX / (r[f"{k}_b200"]["median_ms"] * 1e-3) / 1e9
        r[f"{k}_speedup"] = r[f"{k}_ref"]["median_ms"] / r[f"{k}_b200"]["median_ms"]
    out[f"config2_b{b}"] = r
    del m, X11, X21, D11, D21, rays, D11h, D21h

# ---------------------------------------------------------------- config 3 / 4: GN rays / calib
def gn_case(N, counts, calib, seed):
    g = synth.make_graph(N=N, counts=counts, seed=seed, device=dev)
    if calib:
        g["Xs"] = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])
    iters = 10

    def run(mod):
        def f(T):
            if calib:
                return mod.gauss_newton_calib(T, g["Xs"], g["Cs"], g["K"], g["ii"], g["jj"], g["idx"], g["valid"],
                                              g["Q"], H, W, -10, 1e-6, 1.0, 10.0, 0.0, 1.5, iters, 0.0)
            return mod.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                         0.003, 10.0, 0.0, 1.5, iters, 0.0)
        return f
    setup = lambda: (g["Twc"].clone(),)
    r = {"N": N, "directed_edges": g["M"], "iters": iters}
    r["ref"] = timed(run(ref), setup, warm=2, reps=5)
    r["b200"] = timed(run(be), setup, warm=3, reps=10)
    Ta, Tb = g["Twc"].clone(), g["Twc"].clone()
    run(ref)(Ta); run(be)(Tb)
    torch.cuda.synchronize()
    r["pose_max_abs_diff"] = float((Ta - Tb).abs().max())
    r["pose_max_abs_update"] = float((Tb - g["Twc"]).abs().max())  # the solve did move the poses
    r["iters_per_s_ref"] = iters / (r["ref"]["median_ms"] * 1e-3)
    r["iters_per_s_b200"] = iters / (r["b200"]["median_ms"] * 1e-3)
    r["speedup"] = r["ref"]["median_ms"] / r["b200"]["median_ms"]
    return r


out["config3_gn_rays"] = gn_case(32, None, False, 1236)
out["config4_gn_calib"] = gn_case(64, {1: None, 2: 62, 5: 59, "random": 16}, True, 1237)

s = json.dumps(out, indent=1)
print(s)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(s + "\n")
