"""Caller-side geometry the graph path needs before gauss_newton_calib (reference
mast3r_slam/geometry.py:37-42,107-123): pixel grid, pinhole back-projection and the
ray constraint X <- z K^-1 [u, v, 1]."""
import torch


def get_pixel_coords(b, img_size, device, dtype):
    h, w = img_size
    v, u = torch.meshgrid(torch.arange(h, device=device, dtype=dtype),
                          torch.arange(w, device=device, dtype=dtype), indexing="ij")
    return torch.stack((u, v), -1)[None].expand(b, h, w, 2)


def backproject(p, z, K):
    x = (p[..., 0:1] - K[0, 2]) / K[0, 0]
    y = (p[..., 1:2] - K[1, 2]) / K[1, 1]
    return torch.cat((x * z, y * z, z), -1)


def constrain_points_to_ray(img_size, Xs, K):
    uv = get_pixel_coords(Xs.shape[0], img_size, Xs.device, Xs.dtype).reshape(*Xs.shape[:-1], 2)
    return backproject(uv, Xs[..., 2:3], K).contiguous()
