"""Headless mirror of the viewer's pointmap drawing (reference mast3r_slam/visualization.py).

The reference `Window` draws every keyframe's pointmap with one of two GLSL programs
(`surfelmap.glsl`, `trianglemap.glsl`; `Window.render_pointmap`, visualization.py:334-356) into
an OpenGL window.  No graphics API exists on the B200 boxes, so `HeadlessWindow` keeps the same
drawing call but its target is a device buffer, filled by the CUDA restatement of the shader
arithmetic (csrc/m3r_render.cu); `read()` / `save_ppm()` give the image.  Everything else the
reference window does (imgui, frustums, lines, the moderngl camera) is out of scope.
"""
import math

import torch

import mast3r_slam_backends as _be
from mast3r_slam.sim3 import Sim3


def hex2rgba(h, alpha=1.0):
    """'#1E2326' -> packed RGBA8 word, R in the low byte (visualization.py:53)."""
    h = h.lstrip("#")
    r, g, b = int(h[0:2], 16), int(h[2:4], 16), int(h[4:6], 16)
    return r | (g << 8) | (b << 16) | (int(round(alpha * 255)) << 24)


def perspective(fov_y_deg, aspect, near, far):
    """OpenGL perspective projection, row-major (what camera.proj_mat.gl_matrix() holds, transposed)."""
    t = 1.0 / math.tan(math.radians(fov_y_deg) / 2.0)
    return torch.tensor([[t / aspect, 0, 0, 0], [0, t, 0, 0],
                         [0, 0, -(far + near) / (far - near), -2.0 * far * near / (far - near)],
                         [0, 0, -1, 0]], dtype=torch.float32)


def look_at(eye, target, up=(0.0, -1.0, 0.0)):
    """World -> GL view matrix (x right, y up, -z forward), row-major.  The default `up` is the
    SLAM world's (-y), like the shaders' "assume y up" comment (surfelmap.glsl:51)."""
    eye = torch.as_tensor(eye, dtype=torch.float64)
    f = torch.as_tensor(target, dtype=torch.float64) - eye
    f = f / f.norm()
    s = torch.linalg.cross(f, torch.as_tensor(up, dtype=torch.float64))
    s = s / s.norm()
    u = torch.linalg.cross(s, f)
    m = torch.eye(4, dtype=torch.float64)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[:3, 3] = -(m[:3, :3] @ eye)
    return m.float()


def view_from_camera_pose(T_WC):
    """GL view matrix of a camera sitting at the (computer-vision convention) pose T_WC:
    diag(1,-1,-1,1) * T_WC^-1 (rigid part)."""
    M = T_WC.matrix().reshape(4, 4).double().cpu()
    R = M[:3, :3]
    s = R.norm(dim=0).mean()
    Rn = R / s
    V = torch.eye(4, dtype=torch.float64)
    V[:3, :3] = Rn.T
    V[:3, 3] = -(Rn.T @ M[:3, 3])
    return (torch.diag(torch.tensor([1.0, -1.0, -1.0, 1.0], dtype=torch.float64)) @ V).float()


class HeadlessWindow:
    """`Window` without a window: the clear colour, programs, culling/depth state and the
    `render_pointmap` call of the reference (visualization.py:53-59,77,100-103,334-356)."""

    def __init__(self, width, height, device="cuda", clear="#1E2326"):
        self.width, self.height = int(width), int(height)
        self.device = torch.device(device)
        self.clear_rgba = hex2rgba(clear)
        self.fb = torch.empty((self.height, self.width), dtype=torch.int64, device=self.device)
        self.pointmap_prog = "surfelmap"          # visualization.py:59
        self.C_conf_threshold = 1.5               # WindowMsg, visualization.py:38
        self.culling = True                       # visualization.py:77
        self.show_normal = False
        self.radius = 0.003
        self.slant_threshold = 0.1
        self.camera = torch.eye(4)                # world -> view (camera.gl_matrix() transposed)
        self.proj = perspective(60.0, self.width / self.height, 0.01, 100.0)
        self.clear()

    def clear(self):
        _be.render_clear(self.fb, self.clear_rgba)

    def _models(self, T_WC):
        if isinstance(T_WC, Sim3):
            M = T_WC.matrix()
        else:
            M = torch.as_tensor(T_WC)
        return M.reshape(-1, 4, 4).to(self.device, torch.float32).contiguous()

    def render_pointmap(self, T_WC, w, h, X, C, img, use_img=True, depth_bias=0.0):
        """Reference signature (T_WC, w, h, ptex, ctex, itex, use_img, depth_bias) with the three
        textures given as CUDA tensors: X (K,h,w,3) or (h,w,3), C (..,h,w), img (..,h,w,3) f32.
        T_WC: Sim3 (K poses) or (K,4,4) model matrices.  Draws K keyframes in one launch."""
        w, h = int(w), int(h)
        X = X.reshape(-1, h, w, 3).contiguous()
        C = C.reshape(-1, h, w).contiguous()
        img = img.reshape(-1, h, w, 3).contiguous()
        opts = _be.render_opts(conf_threshold=self.C_conf_threshold, use_img=int(bool(use_img)),
                               depth_bias=float(depth_bias), radius=self.radius,
                               slant_threshold=self.slant_threshold, show_normal=int(self.show_normal),
                               cull_face=int(self.culling))
        _be.render_pointmaps(self.pointmap_prog, X, C, img, self._models(T_WC), self.camera, self.proj,
                             self.fb, opts)

    def read(self):
        """-> (rgba (height,width,4) u8, depth (height,width) f32), top row first."""
        return _be.render_resolve(self.fb)

    def save_ppm(self, path):
        rgba, _ = self.read()
        rgb = rgba[..., :3].contiguous().cpu().numpy()
        with open(path, "wb") as f:
            f.write(b"P6\n%d %d\n255\n" % (self.width, self.height))
            f.write(rgb.tobytes())
