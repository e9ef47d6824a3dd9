"""CPU restatement of matching.match_iterative_proj — TEST INFRASTRUCTURE ONLY.

Follows the reference's call sequence (paths relative to /root/reference/mast3r_slam):
  prep_for_iter_proj   matching.py:25-50  + image.img_gradient image.py:5-38   (torch on CPU)
  iter_proj            backend/src/matching_kernels.cu:119-275                  (oracle C port)
  occlusion test       matching.py:68-76                                        (torch on CPU)
  refine_matches       backend/src/matching_kernels.cu:25-81, fp16              (oracle C port)
  pixel_to_lin         matching.py:13-15
The two kernels have no CPU implementation in the reference; oracle/m3r_oracle.c restates them.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import oracle as orc

CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=0.1, radius=3,
           dilation_max=5)  # config/base.yaml:8-14


def img_gradient(img):  # image.py:5-38
    c = img.shape[1]
    gxk = (1.0 / 32.0) * torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]],
                                      dtype=img.dtype).repeat(c, 1, 1, 1)
    gyk = (1.0 / 32.0) * torch.tensor([[-3.0, -10.0, -3.0], [0.0, 0.0, 0.0], [3.0, 10.0, 3.0]],
                                      dtype=img.dtype).repeat(c, 1, 1, 1)
    pad = F.pad(img, (1, 1, 1, 1), mode="reflect")
    return F.conv2d(pad, gxk, groups=c), F.conv2d(pad, gyk, groups=c)


def prep_for_iter_proj(X11, X21, idx_init=None):  # matching.py:25-50
    b, h, w, _ = X11.shape
    rays = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    img = torch.cat((rays, gx, gy), 1).permute(0, 2, 3, 1).contiguous()
    pts = F.normalize(X21.reshape(b, -1, 3), dim=-1)
    if idx_init is None:
        idx_init = torch.arange(h * w)[None].repeat(b, 1)
    p_init = torch.stack((idx_init % w, idx_init // w), -1).float()
    return img, pts, p_init


def match_iterative_proj(X11, X21, D11, D21, idx_init=None, cfg=CFG, want_p=False):
    """CPU torch tensors in, (idx_1_to_2 (b,n) i64, valid (b,n,1) bool) out."""
    b, h, w = X21.shape[:3]
    img, pts, p_init = prep_for_iter_proj(X11, X21, idx_init)
    p_new, conv = orc.iter_proj(img.numpy(), pts.numpy(), p_init.numpy(), cfg["max_iter"],
                                cfg["lambda_init"], cfg["convergence_thresh"])
    p1 = torch.from_numpy(p_new).long()
    bi = torch.arange(b)[:, None].repeat(1, h * w)
    d = torch.linalg.norm(X11[bi, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - X21, dim=-1)
    valid = torch.from_numpy(conv) & (d < cfg["dist_thresh"]).view(b, -1)
    if cfg["radius"] > 0:
        D11h = D11.half().numpy().view(np.uint16)
        D21h = D21.reshape(b, h * w, -1).half().numpy().view(np.uint16)
        p1 = torch.from_numpy(orc.refine_matches(D11h, D21h, p1.numpy(), cfg["radius"],
                                                 cfg["dilation_max"]))
    idx = p1[..., 0] + w * p1[..., 1]
    if want_p:
        return idx, valid.unsqueeze(-1), torch.from_numpy(p_new)
    return idx, valid.unsqueeze(-1)
