"""Quick per-kernel timing of the fused matching path at b=8 (CUDA events, 20 reps) + equality of the
fused refine against the plain fp16 entry point.  Usage (GPU box): python profiles/time_match.py [b]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

b = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda", 0)
H, W = 384, 512
m = synth.make_match_inputs(b, H, W, 24, seed=1235, device=dev)
X11, X21, D11, D21 = m["X11"], m["X21"], m["D11"], m["D21"]


def t(fn, reps=20):
    for _ in range(3):
        fn()
    a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(reps):
        fn()
    z.record(); torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


rays, D11h = be.match_prep(X11, D11)
p1, valid = be.match_iter_proj_fused(rays, X11, X21, None, 10, 1e-8, 1e-6, 0.1)
idx = be.match_refine_fused(D11h, D21, p1, 3, 5)
(ref,) = be.refine_matches(D11.half(), D21.view(b, H * W, 24).half(), p1, 3, 5)
ok = torch.equal(idx, ref[..., 0] + W * ref[..., 1])
print(f"b={b} prep {t(lambda: be.match_prep(X11, D11)):.4f} ms  iter_proj {t(lambda: be.match_iter_proj_fused(rays, X11, X21, None, 10, 1e-8, 1e-6, 0.1)):.4f} ms  "
      f"refine_fused {t(lambda: be.match_refine_fused(D11h, D21, p1, 3, 5)):.4f} ms  "
      f"refine_f16 {t(lambda: be.refine_matches(D11.half(), D21.view(b, H * W, 24).half(), p1, 3, 5)):.4f} ms  fused==plain {ok}")
