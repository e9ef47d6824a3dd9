"""numpy front-end of oracle/m3r_render_oracle.c — CPU restatement of the viewer's pointmap
shaders (surfelmap.glsl / trianglemap.glsl).  TEST INFRASTRUCTURE ONLY; parity unpinned (no
OpenGL on this box, no golden images in the reference) — see the C file's header.
"""
import ctypes as C

import numpy as np

from . import oracle as _orc


class View(C.Structure):
    _fields_ = [("camera", C.c_float * 16), ("proj", C.c_float * 16), ("width", C.c_int), ("height", C.c_int)]


class Opts(C.Structure):
    _fields_ = [("conf_threshold", C.c_float), ("radius", C.c_float), ("slant_threshold", C.c_float),
                ("depth_bias", C.c_float), ("lightpos", C.c_float * 3), ("phong", C.c_float * 3),
                ("spec", C.c_float), ("base_color", C.c_float * 3), ("use_img", C.c_int),
                ("show_normal", C.c_int), ("cull_face", C.c_int)]


def default_opts(**kw):
    """Uniform defaults of the two programs (surfelmap.glsl:17,99-104; trianglemap.glsl:17-18,111-116)
    and of the viewer state (visualization.py:38,77)."""
    o = Opts()
    o.conf_threshold, o.radius, o.slant_threshold, o.depth_bias = 1.5, 0.003, 0.1, 0.0
    o.lightpos[:] = [0.1, 0.1, 0.0]
    o.phong[:] = [0.3, 0.5, 0.4]
    o.spec = 32.0
    o.base_color[:] = [1.0, 1.0, 1.0]
    o.use_img, o.show_normal, o.cull_face = 1, 0, 1
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(k)
        if k in ("lightpos", "phong", "base_color"):
            getattr(o, k)[:] = list(v)
        else:
            setattr(o, k, v)
    return o


def make_view(camera, proj, width, height):
    v = View()
    v.camera[:] = np.asarray(camera, np.float32).reshape(16).tolist()
    v.proj[:] = np.asarray(proj, np.float32).reshape(16).tolist()
    v.width, v.height = int(width), int(height)
    return v


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def render(program, X, Cf, img, models, camera, proj, width, height, clear_rgba=0xFF26231E, fb=None, **opts):
    """program: "surfelmap" | "trianglemap".  X (K,H,W,3), Cf (K,H,W), img (K,H,W,3), models (K,4,4)
    row-major.  Returns the packed framebuffer (height,width) u64, row 0 = bottom row."""
    lib = _orc.lib()
    X, Cf, img, models = _f32(X), _f32(Cf), _f32(img), _f32(models)
    K, H, W = X.shape[0], X.shape[1], X.shape[2]
    if fb is None:
        fb = np.empty((height, width), np.uint64)
        lib.orc_render_clear(fb.ctypes.data_as(C.c_void_p), int(width), int(height), C.c_uint32(clear_rgba))
    view = make_view(camera, proj, width, height)
    o = default_opts(**opts)
    lib.orc_render(0 if program == "surfelmap" else 1, X.ctypes.data_as(C.c_void_p),
                   Cf.ctypes.data_as(C.c_void_p), img.ctypes.data_as(C.c_void_p),
                   models.ctypes.data_as(C.c_void_p), int(K), int(H), int(W), C.byref(view), C.byref(o),
                   fb.ctypes.data_as(C.c_void_p))
    return fb


def resolve(fb):
    """-> (rgba (h,w,4) u8 top-down, depth (h,w) f32 top-down)"""
    lib = _orc.lib()
    h, w = fb.shape
    rgba = np.empty((h, w, 4), np.uint8)
    depth = np.empty((h, w), np.float32)
    lib.orc_render_resolve(fb.ctypes.data_as(C.c_void_p), int(w), int(h), rgba.ctypes.data_as(C.c_void_p),
                           depth.ctypes.data_as(C.c_void_p))
    return rgba, depth
