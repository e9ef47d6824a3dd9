"""Generate golden vectors from the UNMODIFIED reference extension (oracle/_ref) on a GPU box.

The reference's CUDA kernels have no CPU path and the reference ships no test vectors (SURVEY
F8), so the pin for the CPU oracle is the reference itself, run once on a B200:

    gpurun -- python tests/golden/gen_ref_golden.py        # writes gpurun_out/golden/ref_*.npz
    cp gpurun_out/golden/ref_*.npz tests/golden/            # commit

Inputs are small seeded synthetic cases (tests/synth.py); each file stores the inputs that are
not cheaply regenerable bit-for-bit across devices together with the reference's outputs.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import build_ref  # noqa: E402
from tests import synth  # noqa: E402


def _np(t):
    return t.detach().cpu().numpy()


def main():
    ref = build_ref.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/mast3r_slam_backends_ref.so not found (build it in the container)")
    dev = "cuda:0"
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)

    # ---- matching: iter_proj + refine_matches (fp16 and fp32), b=2, 40x56
    import torch.nn.functional as F

    b, H, W, Fd = 2, 40, 56, 24
    m = synth.make_match_inputs(b, H, W, Fd, seed=4321, device=dev)
    rays = F.normalize(m["X11"], dim=-1).permute(0, 3, 1, 2)
    k = (1.0 / 32.0) * torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]], device=dev)
    pad = F.pad(rays, (1, 1, 1, 1), mode="reflect")
    gx = F.conv2d(pad, k.repeat(3, 1, 1, 1), groups=3)
    gy = F.conv2d(pad, k.t().contiguous().repeat(3, 1, 1, 1), groups=3)
    img = torch.cat((rays, gx, gy), 1).permute(0, 2, 3, 1).contiguous()
    pts = F.normalize(m["X21"].view(b, -1, 3), dim=-1).contiguous()
    lin = torch.arange(H * W, device=dev)[None].repeat(b, 1)
    p0 = torch.stack((lin % W, lin // W), -1).float().contiguous()
    p_new, conv = ref.iter_proj(img, pts, p0, 10, 1e-8, 1e-6)
    torch.cuda.synchronize()
    p1 = p_new.long().contiguous()
    D11h = m["D11"].half().contiguous()
    D21h = m["D21"].view(b, H * W, Fd).half().contiguous()
    (r16,) = ref.refine_matches(D11h, D21h, p1, 3, 5)
    (r32,) = ref.refine_matches(m["D11"].contiguous(), m["D21"].view(b, H * W, Fd).contiguous(), p1, 3, 5)
    (r16_r2,) = ref.refine_matches(D11h, D21h, p1, 2, 3)
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(out_dir, "ref_matching.npz"),
                        img=_np(img), pts=_np(pts), p0=_np(p0), p_new=_np(p_new), conv=_np(conv),
                        D11h=_np(D11h).view(np.uint16), D21h=_np(D21h).view(np.uint16), p1=_np(p1),
                        refine_f16=_np(r16), refine_f32=_np(r32), refine_f16_r2d3=_np(r16_r2))

    # ---- Gauss-Newton: rays / calib / points on a 5-keyframe graph, 24x32
    H, W = 24, 32
    g = synth.make_graph(N=5, counts={1: None, 2: 2}, H=H, W=W, seed=99, device=dev,
                         kf_ids=[2, 3, 5, 8, 13])
    Xc = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])
    store = dict(Twc=_np(g["Twc"]), Xs=_np(g["Xs"]), Xs_calib=_np(Xc), Cs=_np(g["Cs"]), ii=_np(g["ii"]),
                 jj=_np(g["jj"]), idx=_np(g["idx"]), valid=_np(g["valid"]), Q=_np(g["Q"]), K=_np(g["K"]),
                 HW=np.array([H, W]))
    for iters in (1, 10):
        T = g["Twc"].clone()
        (dx,) = ref.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                      0.003, 10.0, 0.0, 1.5, iters, 0.0)
        torch.cuda.synchronize()
        store[f"rays_T_{iters}"], store[f"rays_dx_{iters}"] = _np(T), _np(dx)
        T = g["Twc"].clone()
        (dx,) = ref.gauss_newton_calib(T, Xc, g["Cs"], g["K"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                       H, W, -10, 1e-6, 1.0, 10.0, 0.0, 1.5, iters, 0.0)
        torch.cuda.synchronize()
        store[f"calib_T_{iters}"], store[f"calib_dx_{iters}"] = _np(T), _np(dx)
        T = g["Twc"].clone()
        (dx,) = ref.gauss_newton_points(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                        0.05, 0.0, 1.5, iters, 0.0)
        torch.cuda.synchronize()
        store[f"points_T_{iters}"], store[f"points_dx_{iters}"] = _np(T), _np(dx)
    np.savez_compressed(os.path.join(out_dir, "ref_gn.npz"), **store)
    for f in sorted(os.listdir(out_dir)):
        print(f, os.path.getsize(os.path.join(out_dir, f)), "bytes")


if __name__ == "__main__":
    main()
