"""matching.match / match_iterative_proj with the reference's signatures (reference
mast3r_slam/matching.py:8-90), running as THREE CUDA launches of libm3r_b200.so:

  m3r_match_prep             normalize + 3x3 gradients + pack  (matching.py:25-38, image.py:5-38)
                             and the fp32->fp16 cast of D11     (matching.py:80)
  m3r_match_iter_proj_fused  F.normalize(X21), p_init, LM loop, .long(), occlusion test
                             (matching.py:40-76 around mast3r_slam_backends.iter_proj)
  m3r_match_refine_fused     D21.half(), dilated argmax, pixel_to_lin (matching.py:78-88)

The reference runs ~15 PyTorch kernels around its two CUDA kernels for the same result."""
import torch

import mast3r_slam_backends as _be
from mast3r_slam.config import config


def pixel_to_lin(p1, w):
    return p1[..., 0] + w * p1[..., 1]


def lin_to_pixel(idx_1_to_2, w):
    return torch.stack((idx_1_to_2 % w, idx_1_to_2 // w), dim=-1)


def prep_for_iter_proj(X11, X21, idx_1_to_2_init):
    """-> (rays_with_grad_img (b,h,w,9), pts3d_norm (b,n,3), p_init (b,n,2) f32), the inputs of
    mast3r_slam_backends.iter_proj.  Kept for callers that use the unfused entry point."""
    b, h, w, _ = X11.shape
    rays, _ = _be.match_prep(X11.contiguous())
    pts3d_norm = torch.nn.functional.normalize(X21.reshape(b, -1, 3), dim=-1)
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    return rays, pts3d_norm, lin_to_pixel(idx_1_to_2_init, w).float()


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None):
    """X11, X21 (b,h,w,3) f32; D11, D21 (b,h,w,F) f32; idx_1_to_2_init (b,h*w) i64 or None
    -> idx_1_to_2 (b,h*w) i64, valid (b,h*w,1) bool."""
    cfg = config["matching"]
    b, h, w = X21.shape[:3]
    X11, X21 = X11.contiguous(), X21.contiguous()
    refine = cfg["radius"] > 0
    # float4-packed private ray image: the LM loop gathers it with 128-bit loads
    rays, D11h = _be.match_prep(X11, D11.contiguous() if refine else None, packed=True)
    if idx_1_to_2_init is not None:
        idx_1_to_2_init = idx_1_to_2_init.contiguous()
    p1, valid = _be.match_iter_proj_fused(rays, X11, X21, idx_1_to_2_init, cfg["max_iter"],
                                          cfg["lambda_init"], cfg["convergence_thresh"],
                                          cfg["dist_thresh"])
    if refine:
        idx_1_to_2 = _be.match_refine_fused(D11h, D21.contiguous(), p1, cfg["radius"],
                                            cfg["dilation_max"])
    else:
        idx_1_to_2 = pixel_to_lin(p1, w)
    return idx_1_to_2, valid.unsqueeze(-1)


def match(X11, X21, D11, D21, idx_1_to_2_init=None):
    return match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init)
