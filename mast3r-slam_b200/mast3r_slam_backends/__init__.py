"""mast3r_slam_backends — B200 (sm_100a) drop-in for MASt3R-SLAM's native extension.

Same import name, callables, positional signatures and return arity as the reference pybind
module (reference mast3r_slam/backend/src/gn.cpp:116-122, include/gn.h):

    iter_proj, refine_matches, gauss_newton_rays, gauss_newton_calib, gauss_newton_points

so `matching.match_iterative_proj` (matching.py:60,79) and `FactorGraph.solve_GN_rays/calib`
(global_opt.py:140,190) call it unchanged.  Everything is executed by hand-written CUDA
kernels in libm3r_b200.so, bound through its C ABI (include/m3r_b200.h) with ctypes; torch is
used only to allocate outputs/workspace and to supply the current CUDA stream.  There is no
CPU path: tensors must be CUDA tensors and a missing library is an ImportError.

Additive entry points (not in the reference): pose_retr, match_prep, match_iter_proj_fused,
match_refine_fused, track_ray_dist, gn_prepare / gn_accumulate / gn_solve_step /
gauss_newton_sharded (edge-sharded multi-GPU solve), gn_blocks_reference_layout (tests).
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# M3R_B200_LIB: load another build of the SAME library (experiments under profiles/, e.g. a
# -DM3R_GN_PRECISE variant); it is still this package's CUDA code, never a fallback
_LIB_PATH = os.environ.get("M3R_B200_LIB") or os.path.join(_HERE, "..", "lib", "libm3r_b200.so")


class _GNParams(C.Structure):
    _fields_ = [("mode", C.c_int), ("sigma_a", C.c_float), ("sigma_b", C.c_float),
                ("C_thresh", C.c_float), ("Q_thresh", C.c_float),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("height", C.c_int), ("width", C.c_int), ("pixel_border", C.c_int),
                ("z_eps", C.c_float), ("K_dev", C.c_void_p)]


MAX_PEERS = 16


class _GNPeers(C.Structure):
    _fields_ = [("blocks", C.c_void_p * MAX_PEERS), ("flags", C.c_void_p * MAX_PEERS),
                ("world", C.c_int), ("rank", C.c_int), ("epoch", C.c_int)]


GN_RAYS, GN_CALIB, GN_POINTS = 0, 1, 2


class _RenderView(C.Structure):
    _fields_ = [("camera", C.c_float * 16), ("proj", C.c_float * 16), ("width", C.c_int),
                ("height", C.c_int)]


class _RenderOpts(C.Structure):
    _fields_ = [("conf_threshold", C.c_float), ("radius", C.c_float), ("slant_threshold", C.c_float),
                ("depth_bias", C.c_float), ("lightpos", C.c_float * 3), ("phong", C.c_float * 3),
                ("spec", C.c_float), ("base_color", C.c_float * 3), ("use_img", C.c_int),
                ("show_normal", C.c_int), ("cull_face", C.c_int)]

EXPORTS = [
    "m3r_version", "m3r_error_string", "m3r_last_cuda_error", "m3r_launch_count",
    "m3r_iter_proj", "m3r_refine_matches_f16", "m3r_refine_matches_f32", "m3r_match_prep",
    "m3r_match_iter_proj_fused", "m3r_match_refine_fused", "m3r_pose_retr",
    "m3r_gn_workspace_bytes", "m3r_gauss_newton", "m3r_gn_prepare", "m3r_gn_accumulate",
    "m3r_gn_solve_step", "m3r_gn_done_flag", "m3r_gn_blocks", "m3r_track_workspace_bytes",
    "m3r_track_ray_dist", "m3r_track_calib", "m3r_debug_chol_trace", "m3r_debug_chol_delay", "m3r_edge_build", "m3r_pointmap_update", "m3r_gn_accumulate_p2p", "m3r_gn_solve_step_p2p",
    "m3r_match_prep_packed", "m3r_match_iter_proj_packed", "m3r_img_gradient",
    "m3r_render_default_opts", "m3r_render_clear", "m3r_render_surfels", "m3r_render_triangles",
    "m3r_render_resolve",
    "m3r_gn_status", "m3r_gn_prepare_two_way", "m3r_gn_accumulate_packed", "m3r_gauss_newton_packed",
    "m3r_edge_build_packed", "m3r_edge_pack", "m3r_edge_unpack",
    "m3r_track_ray_dist_noalloc", "m3r_debug_chol64_tasklist",
]


def _load():
    path = os.path.abspath(_LIB_PATH)
    if not os.path.exists(path):
        raise ImportError(
            f"mast3r_slam_backends: {path} not found. Build it with "
            "`python mast3r-slam_b200/build.py` (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(path)
    vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    lib.m3r_error_string.restype = C.c_char_p
    lib.m3r_error_string.argtypes = [i32]
    lib.m3r_launch_count.restype = C.c_ulonglong
    lib.m3r_iter_proj.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, f32, f32, vp, vp, vp]
    lib.m3r_refine_matches_f16.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp]
    lib.m3r_refine_matches_f32.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp]
    lib.m3r_match_prep.argtypes = [vp, vp, i32, i32, i32, i32, vp, vp, vp]
    lib.m3r_match_iter_proj_fused.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, f32,
                                              vp, vp, vp, vp]
    lib.m3r_match_prep_packed.argtypes = lib.m3r_match_prep.argtypes
    lib.m3r_img_gradient.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    lib.m3r_match_iter_proj_packed.argtypes = lib.m3r_match_iter_proj_fused.argtypes
    lib.m3r_match_refine_fused.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]
    lib.m3r_pose_retr.argtypes = [vp, vp, i32, i32, vp]
    lib.m3r_gn_workspace_bytes.restype = sz
    lib.m3r_gn_workspace_bytes.argtypes = [i32, i32, i32, i32]
    gp = C.POINTER(_GNParams)
    lib.m3r_gauss_newton.argtypes = [gp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32,
                                     vp, vp, sz, vp]
    lib.m3r_gn_prepare.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, vp, sz, vp]
    lib.m3r_gn_accumulate.argtypes = [gp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, sz, vp]
    lib.m3r_gn_solve_step.argtypes = [vp, vp, i32, i32, f32, vp, vp, sz, vp]
    lib.m3r_gn_done_flag.restype = vp
    lib.m3r_gn_done_flag.argtypes = [vp]
    lib.m3r_gn_blocks.restype = vp
    lib.m3r_gn_blocks.argtypes = [vp, i32, i32, i32, i32]
    lib.m3r_debug_chol_trace.argtypes = [vp, i32]
    lib.m3r_debug_chol_trace.restype = None
    lib.m3r_debug_chol_delay.argtypes = [i32, i32]
    lib.m3r_debug_chol_delay.restype = None
    lib.m3r_track_workspace_bytes.restype = sz
    lib.m3r_track_workspace_bytes.argtypes = [i32]
    lib.m3r_track_ray_dist.argtypes = [vp, vp, vp, vp, i32, f32, f32, f32, i32, f32, f32, vp, vp,
                                       vp, vp, sz, vp]
    lib.m3r_debug_chol64_tasklist.argtypes = [i32, vp, i32]
    lib.m3r_track_ray_dist_noalloc.argtypes = [vp, vp, vp, vp, i32, f32, f32, f32, i32, f32, f32, vp, vp, vp, vp]
    lib.m3r_track_calib.argtypes = [vp, vp, vp, vp, vp, i32, vp, i32, i32, i32, f32,
                                    f32, f32, f32, i32, f32, f32, vp, vp, vp, vp, sz, vp]
    pp = C.POINTER(_GNPeers)
    lib.m3r_gn_accumulate_p2p.argtypes = [gp, vp, vp, vp, vp, i32, i32, i32, i32, i32, pp, vp, sz, vp]
    lib.m3r_gn_solve_step_p2p.argtypes = [vp, pp, i32, i32, f32, vp, vp, sz, vp]
    lib.m3r_edge_build.argtypes = [vp, vp, vp, vp, i32, i32, f32, vp, vp, vp]
    lib.m3r_edge_build_packed.argtypes = [vp, vp, vp, vp, i32, i32, f32, vp, vp, vp, vp]
    lib.m3r_edge_pack.argtypes = [vp, vp, vp, sz, vp, vp]
    lib.m3r_edge_unpack.argtypes = [vp, sz, vp, vp, vp, vp]
    lib.m3r_gn_status.restype = vp
    lib.m3r_gn_status.argtypes = [vp]
    lib.m3r_gn_prepare_two_way.argtypes = lib.m3r_gn_prepare.argtypes
    lib.m3r_gn_accumulate_packed.argtypes = [gp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, pp, vp, sz, vp]
    lib.m3r_gauss_newton_packed.argtypes = [gp, vp, vp, vp, vp, vp, i32, vp, vp, i32, i32, i32, i32, i32, f32,
                                            vp, vp, sz, vp]
    lib.m3r_pointmap_update.argtypes = [vp, vp, vp, vp, vp, i32, i32, vp]
    rv, ro = C.POINTER(_RenderView), C.POINTER(_RenderOpts)
    lib.m3r_render_default_opts.restype = None
    lib.m3r_render_default_opts.argtypes = [ro]
    lib.m3r_render_clear.argtypes = [vp, i32, i32, C.c_uint32, vp]
    lib.m3r_render_surfels.argtypes = [vp, vp, vp, vp, i32, i32, i32, rv, ro, vp, vp]
    lib.m3r_render_triangles.argtypes = [vp, vp, vp, vp, i32, i32, i32, rv, ro, vp, vp]
    lib.m3r_render_resolve.argtypes = [vp, i32, i32, vp, vp, vp]
    return lib


_lib = _load()
LIBRARY_PATH = os.path.abspath(_LIB_PATH)


def launch_count():
    """Kernels launched by libm3r_b200.so in this process (bench `gpu_launches`)."""
    return int(_lib.m3r_launch_count())


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what}: {_lib.m3r_error_string(rc).decode()} "
                           f"(code {rc}, cudaError {_lib.m3r_last_cuda_error()})")


def _stream(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _req(t, name, dtype=None, ndim=None):
    """Reference checks contiguity only (gn.h:5); we also insist on CUDA + dtype + rank."""
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f"{name} must be a torch.Tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (mast3r_slam_backends has no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    if ndim is not None and t.dim() != ndim:
        raise RuntimeError(f"{name} must have {ndim} dimensions, got {t.dim()}")
    return t


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


# --------------------------------------------------------------------------- matching

def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    """-> [p_new (b,n,2) f32, converged (b,n) bool]   (gn.h:89-95)"""
    img = _req(rays_img_with_grad, "rays_img_with_grad", torch.float32, 4)
    pts = _req(pts_3d_norm, "pts_3d_norm", torch.float32, 3)
    p0 = _req(p_init, "p_init", torch.float32, 3)
    b, H, W, c = img.shape
    if c != 9:
        raise RuntimeError("rays_img_with_grad must have 9 channels")
    n = p0.shape[1]
    with torch.cuda.device(img.device):
        p_new = torch.empty((b, n, 2), dtype=torch.float32, device=img.device)
        conv = torch.empty((b, n), dtype=torch.bool, device=img.device)
        _check(_lib.m3r_iter_proj(_ptr(img), _ptr(pts), _ptr(p0), b, H, W, n, int(max_iter),
                                  float(lambda_init), float(cost_thresh), _ptr(p_new), _ptr(conv),
                                  _stream(img)), "iter_proj")
    return [p_new, conv]


def refine_matches(D11, D21, p1, radius, dilation_max):
    """-> [p1_new (b,n,2) i64]   (gn.h:112-117); D11/D21 float16 or float32."""
    D11 = _req(D11, "D11", None, 4)
    D21 = _req(D21, "D21", D11.dtype, 3)
    p1 = _req(p1, "p1", torch.int64, 3)
    b, H, W, F = D11.shape
    n = p1.shape[1]
    with torch.cuda.device(D11.device):
        out = torch.empty_like(p1)
        if D11.dtype == torch.float16:
            fn = _lib.m3r_refine_matches_f16
        elif D11.dtype == torch.float32:
            fn = _lib.m3r_refine_matches_f32
        else:
            raise RuntimeError("refine_matches: D11 must be float16 or float32")
        _check(fn(_ptr(D11), _ptr(D21), _ptr(p1), b, H, W, n, F, int(radius), int(dilation_max),
                  _ptr(out), _stream(D11)), "refine_matches")
    return [out]


def match_prep(X11, D11=None, packed=False):
    """Fused prep_for_iter_proj (matching.py:25-50) + D11.half() (matching.py:80).
    -> (rays_img_with_grad (b,H,W,9) f32, D11h (b,H,W,F) f16 or None); packed=True gives the
    private (b,H,W,12) float4 layout [ray,0 | d/du,0 | d/dv,0] that match_iter_proj_fused gathers
    with 128-bit loads."""
    X11 = _req(X11, "X11", torch.float32, 4)
    b, H, W, _ = X11.shape
    with torch.cuda.device(X11.device):
        rays = torch.empty((b, H, W, 12 if packed else 9), dtype=torch.float32, device=X11.device)
        D11h, F = None, 0
        if D11 is not None:
            D11 = _req(D11, "D11", torch.float32, 4)
            F = D11.shape[-1]
            D11h = torch.empty(D11.shape, dtype=torch.float16, device=X11.device)
        fn = _lib.m3r_match_prep_packed if packed else _lib.m3r_match_prep
        _check(fn(_ptr(X11), _ptr(D11), b, H, W, F, _ptr(rays), _ptr(D11h), _stream(X11)), "match_prep")
    return rays, D11h


def img_gradient(img):
    """image.img_gradient (image.py:5-38): img (b,c,h,w) f32 -> (gx, gy) of the same shape."""
    img = _req(img, "img", torch.float32, 4)
    b, c, h, w = img.shape
    with torch.cuda.device(img.device):
        gx = torch.empty_like(img)
        gy = torch.empty_like(img)
        _check(_lib.m3r_img_gradient(_ptr(img), b * c, h, w, _ptr(gx), _ptr(gy), _stream(img)), "img_gradient")
    return gx, gy


def match_iter_proj_fused(rays_img_with_grad, X11, X21, idx_init, max_iter, lambda_init,
                          cost_thresh, dist_thresh, want_p_new=False):
    """-> (p1 (b,n,2) i64, valid (b,n) bool[, p_new (b,n,2) f32])"""
    img = _req(rays_img_with_grad, "rays_img_with_grad", torch.float32, 4)
    X11 = _req(X11, "X11", torch.float32, 4)
    X21 = _req(X21, "X21", torch.float32, 4)
    b, H, W, ch = img.shape
    if ch not in (9, 12):
        raise RuntimeError("rays_img_with_grad must have 9 channels (reference layout) or 12 (packed)")
    n = H * W
    if idx_init is not None:
        idx_init = _req(idx_init, "idx_1_to_2_init", torch.int64, 2)
    with torch.cuda.device(img.device):
        p1 = torch.empty((b, n, 2), dtype=torch.int64, device=img.device)
        valid = torch.empty((b, n), dtype=torch.bool, device=img.device)
        p_new = torch.empty((b, n, 2), dtype=torch.float32, device=img.device) if want_p_new else None
        fn = _lib.m3r_match_iter_proj_packed if ch == 12 else _lib.m3r_match_iter_proj_fused
        _check(fn(
            _ptr(img), _ptr(X11), _ptr(X21), _ptr(idx_init), b, H, W, int(max_iter),
            float(lambda_init), float(cost_thresh), float(dist_thresh), _ptr(p1), _ptr(valid),
            _ptr(p_new), _stream(img)), "match_iter_proj_fused")
    return (p1, valid, p_new) if want_p_new else (p1, valid)


def match_refine_fused(D11h, D21, p1, radius, dilation_max):
    """-> idx_1_to_2 (b,n) i64"""
    D11h = _req(D11h, "D11h", torch.float16, 4)
    D21 = _req(D21, "D21", torch.float32)
    p1 = _req(p1, "p1", torch.int64, 3)
    b, H, W, F = D11h.shape
    with torch.cuda.device(D11h.device):
        idx = torch.empty((b, H * W), dtype=torch.int64, device=D11h.device)
        _check(_lib.m3r_match_refine_fused(_ptr(D11h), _ptr(D21), _ptr(p1), b, H, W, F,
                                           int(radius), int(dilation_max), _ptr(idx),
                                           _stream(D11h)), "match_refine_fused")
    return idx


# --------------------------------------------------------------------------- Sim3 / GN

def pose_retr(poses, dx, num_fix=1):
    """In place: T_k <- Exp(dx_{k-num_fix}) * T_k   (pose_retr_kernel, gn_kernels.cu:415-453)"""
    poses = _req(poses, "poses", torch.float32, 2)
    dx = _req(dx, "dx", torch.float32, 2)
    _check(_lib.m3r_pose_retr(_ptr(poses), _ptr(dx), poses.shape[0], int(num_fix), _stream(poses)),
           "pose_retr")
    return poses


def _params(mode, sigma_a, sigma_b, C_thresh, Q_thresh, K=None, height=0, width=0,
            pixel_border=0, z_eps=0.0, device=None):
    p = _GNParams()
    p.mode = mode
    p.sigma_a, p.sigma_b = float(sigma_a), float(sigma_b)
    p.C_thresh, p.Q_thresh = float(C_thresh), float(Q_thresh)
    if K is not None:
        # the kernels read fx, fy, cx, cy from the DEVICE matrix, like the reference
        # (gn_kernels.cu:1274-1279): no .to("cpu"), hence no hidden host synchronisation
        Kd = K.detach()
        if not (Kd.is_cuda and Kd.dtype == torch.float32 and Kd.is_contiguous() and Kd.shape == (3, 3)):
            if not Kd.is_cuda and device is None:
                raise RuntimeError("K must be a CUDA tensor (or pass the target device)")
            Kd = Kd.to(device=device if not Kd.is_cuda else Kd.device, dtype=torch.float32).contiguous()
            if Kd.shape != (3, 3):
                raise RuntimeError("K must be (3,3)")
        p._K_keepalive = Kd  # the struct holds a raw pointer
        p.K_dev = Kd.data_ptr()
    p.height, p.width, p.pixel_border, p.z_eps = int(height), int(width), int(pixel_border), float(z_eps)
    return p


def _gn_inputs(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q):
    Twc = _req(Twc, "Twc", torch.float32, 2)
    Xs = _req(Xs, "Xs", torch.float32, 3)
    Cs = _req(Cs, "Cs", torch.float32)
    ii = _req(ii, "ii", torch.int64, 1)
    jj = _req(jj, "jj", torch.int64, 1)
    idx = _req(idx_ii2jj, "idx_ii2jj", torch.int64, 2)
    vm = _req(valid_match, "valid_match", torch.bool)
    Q = _req(Q, "Q", torch.float32)
    N, n = Xs.shape[0], Xs.shape[1]
    if Twc.shape[0] != N or Twc.shape[1] != 8:
        raise RuntimeError("Twc must be (num_poses, 8)")
    return Twc, Xs, Cs, ii, jj, idx, vm, Q, N, n


def _gauss_newton(prm, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh):
    Twc, Xs, Cs, ii, jj, idx, vm, Q, N, n = _gn_inputs(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q)
    M = ii.shape[0]
    with torch.cuda.device(Twc.device):
        dx = torch.empty((N - 1, 7), dtype=torch.float32, device=Twc.device)
        nbytes = _lib.m3r_gn_workspace_bytes(N, M, M, n)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=Twc.device)
        _check(_lib.m3r_gauss_newton(C.byref(prm), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj),
                                     _ptr(idx), _ptr(vm), _ptr(Q), N, n, M, int(max_iter),
                                     float(delta_thresh), _ptr(dx), _ptr(ws), nbytes, _stream(Twc)),
               "gauss_newton")
    return [dx]


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist,
                      C_thresh, Q_thresh, max_iter, delta_thresh):
    """-> [dx (N-1,7)]; Twc updated in place   (gn.h:45-55)"""
    prm = _params(GN_RAYS, sigma_ray, sigma_dist, C_thresh, Q_thresh)
    return _gauss_newton(prm, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width,
                       pixel_border, z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter,
                       delta_thresh):
    """-> [dx (N-1,7)]; Twc updated in place   (gn.h:74-87)"""
    prm = _params(GN_CALIB, sigma_pixel, sigma_depth, C_thresh, Q_thresh, K, height, width,
                  pixel_border, z_eps, device=Twc.device)
    return _gauss_newton(prm, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh,
                        Q_thresh, max_iter, delta_thresh):
    """-> [dx (N-1,7)]; Twc updated in place   (gn.h:22-31)"""
    prm = _params(GN_POINTS, sigma_point, 0.0, C_thresh, Q_thresh)
    return _gauss_newton(prm, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh)


def _req_packed(t, name, n=None):
    """Packed edge records (include/m3r_b200.h m3r_packed_match): (E,n) int64, one 8-byte record
    per correspondence (low word idx i32, high word the sign-folded Q f32)."""
    t = _req(t, name, torch.int64, 2)
    if n is not None and t.shape[1] != n:
        raise RuntimeError(f"{name} must have {n} records per edge, got {t.shape[1]}")
    if t.data_ptr() % 8:
        raise RuntimeError(f"{name} must be 8-byte aligned")
    return t


def gauss_newton_packed(prm, Twc, Xs, Cs, ii, jj, pk0, pk1, max_iter, delta_thresh, two_way=False):
    """The gauss_newton_* loop over PACKED edges in up to two row blocks (m3r_gauss_newton_packed).
    two_way=True: ii/jj are the E undirected edges, pk0 their forward and pk1 their reverse
    records — FactorGraph.solve_GN_* without prep_two_way_edges' concatenations.
    two_way=False: ii/jj are the M directed edges, rows [0,len(pk0)) in pk0, the rest in pk1 (may be None)."""
    Twc = _req(Twc, "Twc", torch.float32, 2)
    Xs = _req(Xs, "Xs", torch.float32, 3)
    Cs = _req(Cs, "Cs", torch.float32)
    ii = _req(ii, "ii", torch.int64, 1)
    jj = _req(jj, "jj", torch.int64, 1)
    N, n = Xs.shape[0], Xs.shape[1]
    if Twc.shape[0] != N or Twc.shape[1] != 8:
        raise RuntimeError("Twc must be (num_poses, 8)")
    pk0 = _req_packed(pk0, "pk0", n)
    pk1 = _req_packed(pk1, "pk1", n) if pk1 is not None else None
    M = ii.shape[0] * (2 if two_way else 1)
    split = pk0.shape[0]
    if split + (pk1.shape[0] if pk1 is not None else 0) != M:
        raise RuntimeError(f"packed edge blocks hold {split + (pk1.shape[0] if pk1 is not None else 0)} "
                           f"edges, the edge list has {M}")
    with torch.cuda.device(Twc.device):
        dx = torch.empty((N - 1, 7), dtype=torch.float32, device=Twc.device)
        nbytes = _lib.m3r_gn_workspace_bytes(N, M, M, n)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=Twc.device)
        _check(_lib.m3r_gauss_newton_packed(C.byref(prm), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj),
                                            1 if two_way else 0, _ptr(pk0), _ptr(pk1), split, N, n, M,
                                            int(max_iter), float(delta_thresh), _ptr(dx), _ptr(ws), nbytes,
                                            _stream(Twc)), "gauss_newton_packed")
    return [dx]


def edge_pack(idx, valid_match, Q, out=None):
    """Reference layout (idx (E,n) i64, valid_match (E,n,1) bool, Q (E,n,1) f32) -> packed (E,n) int64."""
    idx = _req(idx, "idx", torch.int64, 2)
    vm = _req(valid_match, "valid_match", torch.bool)
    Q = _req(Q, "Q", torch.float32)
    if vm.numel() != idx.numel() or Q.numel() != idx.numel():
        raise RuntimeError("edge_pack: idx, valid_match and Q must have the same number of points")
    with torch.cuda.device(idx.device):
        if out is None:
            out = torch.empty(idx.shape, dtype=torch.int64, device=idx.device)
        else:
            out = _req_packed(out, "out", idx.shape[1])
            if out.shape != idx.shape:
                raise RuntimeError("edge_pack: out has the wrong shape")
        _check(_lib.m3r_edge_pack(_ptr(idx), _ptr(vm), _ptr(Q), idx.numel(), _ptr(out), _stream(idx)), "edge_pack")
    return out


def edge_unpack(pk):
    """packed (E,n) int64 -> (idx (E,n) i64, valid_match (E,n,1) bool, Q (E,n,1) f32), lossless."""
    pk = _req_packed(pk, "pk")
    E, n = pk.shape
    with torch.cuda.device(pk.device):
        idx = torch.empty((E, n), dtype=torch.int64, device=pk.device)
        vm = torch.empty((E, n, 1), dtype=torch.bool, device=pk.device)
        Q = torch.empty((E, n, 1), dtype=torch.float32, device=pk.device)
        _check(_lib.m3r_edge_unpack(_ptr(pk), pk.numel(), _ptr(idx), _ptr(vm), _ptr(Q), _stream(pk)), "edge_unpack")
    return idx, vm, Q


class GNSolver:
    """Stepwise Gauss-Newton over a slice [edge_begin, edge_begin+edge_count) of the directed edge
    list: prepare once, then per iteration accumulate() -> (exchange blocks) -> solve_step().
    `ii`/`jj` are the FULL edge lists; the per-edge streams are the LOCAL slice, either in the
    reference layout (idx_local, valid_local, Q_local) or — idx_local an (Ml,n) int64 tensor of
    packed records and valid_local = Q_local = None — in the packed layout (edge_pack /
    edge_build_packed), optionally continued in a second row block `packed_tail`."""

    def __init__(self, prm, Twc, Xs, Cs, ii, jj, idx_local, valid_local, Q_local, edge_begin=0,
                 peers=None, packed_tail=None):
        self.packed = valid_local is None and Q_local is None
        if self.packed:
            self.Twc = _req(Twc, "Twc", torch.float32, 2)
            self.Xs = _req(Xs, "Xs", torch.float32, 3)
            self.Cs = _req(Cs, "Cs", torch.float32)
            self.ii = _req(ii, "ii", torch.int64, 1)
            self.jj = _req(jj, "jj", torch.int64, 1)
            self.N, self.n = self.Xs.shape[0], self.Xs.shape[1]
            self.idx = _req_packed(idx_local, "packed", self.n)
            self.tail = _req_packed(packed_tail, "packed_tail", self.n) if packed_tail is not None else None
            self.vm = self.Q = None
            self.Ml = self.idx.shape[0] + (self.tail.shape[0] if self.tail is not None else 0)
        else:
            (self.Twc, self.Xs, self.Cs, self.ii, self.jj, self.idx, self.vm, self.Q, self.N,
             self.n) = _gn_inputs(Twc, Xs, Cs, ii, jj, idx_local, valid_local, Q_local)
            self.tail = None
            self.Ml = self.idx.shape[0]
        self.prm = prm
        self.M = self.ii.shape[0]
        self.edge_begin = int(edge_begin)
        dev = Twc.device
        with torch.cuda.device(dev):
            self.nbytes = _lib.m3r_gn_workspace_bytes(self.N, self.M, self.Ml, self.n)
            self.ws = torch.empty((self.nbytes,), dtype=torch.uint8, device=dev)
            self.peers = peers  # PeerBlocks: exchange fused into the kernels over NVLink peer memory
            self.blocks = peers.blocks if peers is not None else torch.zeros((self.M, 56), dtype=torch.float64,
                                                                              device=dev)
            self.dx = torch.zeros((self.N - 1, 7), dtype=torch.float32, device=dev)
            _check(_lib.m3r_gn_prepare(_ptr(self.Xs), _ptr(self.Cs), _ptr(self.ii), _ptr(self.jj),
                                       self.N, self.n, self.M, self.Ml, _ptr(self.ws), self.nbytes,
                                       _stream(Twc)), "gn_prepare")

    def accumulate(self):
        peers = C.byref(self.peers.next_epoch()) if self.peers is not None else None
        with torch.cuda.device(self.Twc.device):
            if self.packed:
                _check(_lib.m3r_gn_accumulate_packed(C.byref(self.prm), _ptr(self.Twc), _ptr(self.idx),
                                                     _ptr(self.tail), self.idx.shape[0], self.N, self.n, self.M,
                                                     self.edge_begin, self.Ml,
                                                     None if self.peers is not None else _ptr(self.blocks), peers,
                                                     _ptr(self.ws), self.nbytes, _stream(self.Twc)),
                       "gn_accumulate_packed")
            elif self.peers is not None:
                _check(_lib.m3r_gn_accumulate_p2p(C.byref(self.prm), _ptr(self.Twc), _ptr(self.idx),
                                                  _ptr(self.vm), _ptr(self.Q), self.N, self.n, self.M,
                                                  self.edge_begin, self.Ml, peers,
                                                  _ptr(self.ws), self.nbytes, _stream(self.Twc)),
                       "gn_accumulate_p2p")
            else:
                _check(_lib.m3r_gn_accumulate(C.byref(self.prm), _ptr(self.Twc), _ptr(self.idx),
                                              _ptr(self.vm), _ptr(self.Q), self.N, self.n, self.M,
                                              self.edge_begin, self.Ml, _ptr(self.blocks), _ptr(self.ws),
                                              self.nbytes, _stream(self.Twc)), "gn_accumulate")
        if self.peers is not None:
            self.blocks = self.peers.blocks
        return self.blocks

    def solve_step(self, delta_thresh):
        if self.peers is not None:
            with torch.cuda.device(self.Twc.device):
                _check(_lib.m3r_gn_solve_step_p2p(_ptr(self.Twc), C.byref(self.peers.struct), self.N, self.M,
                                                  float(delta_thresh), _ptr(self.dx), _ptr(self.ws),
                                                  self.nbytes, _stream(self.Twc)), "gn_solve_step_p2p")
            return self.dx
        with torch.cuda.device(self.Twc.device):
            _check(_lib.m3r_gn_solve_step(_ptr(self.Twc), _ptr(self.blocks), self.N, self.M,
                                          float(delta_thresh), _ptr(self.dx), _ptr(self.ws),
                                          self.nbytes, _stream(self.Twc)), "gn_solve_step")
        return self.dx

    def local_blocks(self):
        return self.blocks[self.edge_begin:self.edge_begin + self.Ml]

    def status(self):
        """Host copy (one synchronising 16-byte read) of the solve's status words (m3r_gn_status):
        dict(done, chol_failed, iterations, peer_timeout)."""
        off = _lib.m3r_gn_status(_ptr(self.ws)) - self.ws.data_ptr()
        w = self.ws[off:off + 16].view(torch.int32).tolist()
        return dict(done=bool(w[0]), chol_failed=bool(w[1]), iterations=w[2], peer_timeout=bool(w[3]))


def gn_blocks_reference_layout(blocks):
    """(M,56) f64 per-edge blocks -> Hs (4,M,7,7), gs (2,M,7) in the reference's layout
    (gn_kernels.cu:1096-1137): Hs[0]=Hs[3]=S, Hs[1]=Hs[2]=-S, gs[0]=-v, gs[1]=v."""
    S = blocks[:, :49].reshape(-1, 7, 7)
    v = blocks[:, 49:56]
    return torch.stack([S, -S, -S, S]), torch.stack([-v, v])


def shard_edges(num_edges, world, rank):
    """Contiguous, balanced slices of the directed edge list -> (begin, count, counts of all ranks).
    Every edge costs the same n points, so balancing by count balances the work."""
    base, rem = divmod(int(num_edges), int(world))
    counts = [base + (1 if r < rem else 0) for r in range(world)]
    begin = sum(counts[:rank])
    return begin, counts[rank], counts


def exchange_blocks(blocks, begin, count, counts, group=None):
    """All-gather the per-edge blocks: rank r owns rows [sum(counts[:r]), +counts[r]) of
    `blocks` (M,56) f64; on return every rank holds all M rows.  One collective per GN iteration
    (M*448 bytes: 90 KB at 2E=200, 3.7 MB at 2E=8192) — NCCL over NVLink on GPUs, gloo in the
    CPU tests.  Because each edge's block is computed by exactly one rank in a fixed order, the
    gathered array — and hence the replicated solve — is bit-identical on all ranks and to the
    single-GPU result (a sum-allreduce of the assembled Hessian would not be)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if world == 1:
        return blocks
    local = blocks[begin:begin + count]
    if len(set(counts)) == 1:
        dist.all_gather_into_tensor(blocks, local.clone(), group=group)
        return blocks
    cap = max(counts)
    send = blocks.new_zeros((cap, blocks.shape[1]))
    send[:count] = local
    recv = blocks.new_empty((world * cap, blocks.shape[1]))
    dist.all_gather_into_tensor(recv, send, group=group)
    off = 0
    for r, c in enumerate(counts):
        blocks[off:off + c] = recv[r * cap:r * cap + c]
        off += c
    return blocks


class PeerBlocks:
    """Per-edge block buffer + arrival flags of every rank, mapped into every process (NVLink peer
    memory via torch.distributed._symmetric_memory).  With it the block exchange of the sharded solve
    needs no collective call: the producing kernel stores into all ranks' buffers and raises a flag,
    the consuming kernel waits for the flags on the device (include/m3r_b200.h, m3r_gn_peers).
    Construction is collective (every rank of `group`, same num_edges); reuse the object across solves."""

    _cache = {}          # (num_edges, device, group) -> PeerBlocks, insertion-ordered
    _cache_max = 4       # distinct graph sizes kept mapped; the oldest mapping is dropped beyond that

    def __init__(self, num_edges, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        group = group or dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > MAX_PEERS:
            raise RuntimeError(f"PeerBlocks supports at most {MAX_PEERS} ranks")
        with torch.cuda.device(device):
            # two buffers, used alternately by epoch parity: a fast rank may already push iteration
            # k+1 while a slow one still assembles iteration k; iteration k+2 cannot start anywhere
            # before every rank has raised its flag for k+1, i.e. has finished reading k
            self.buffers = symm.empty((2, int(num_edges), 56), dtype=torch.float64, device=device)
            self.blocks = self.buffers[0]
            self.flags = symm.empty((64,), dtype=torch.int32, device=device)
            self.buffers.zero_(); self.flags.zero_()
            torch.cuda.synchronize(device)
            hb, hf = symm.rendezvous(self.buffers, group), symm.rendezvous(self.flags, group)
            hf.barrier()
        self._handles = (hb, hf)
        self.struct = _GNPeers()
        self._base = [int(hb.buffer_ptrs[r]) for r in range(self.world)]
        self._stride = int(num_edges) * 56 * 8
        for r in range(self.world):
            self.struct.blocks[r] = self._base[r]
            self.struct.flags[r] = hf.buffer_ptrs[r]
        self.struct.world, self.struct.rank, self.struct.epoch = self.world, self.rank, 0

    def next_epoch(self):
        self.struct.epoch += 1
        par = self.struct.epoch & 1
        for r in range(self.world):
            self.struct.blocks[r] = self._base[r] + par * self._stride
        self.blocks = self.buffers[par]
        return self.struct

    @classmethod
    def get(cls, num_edges, device, group=None):
        key = (int(num_edges), str(device), id(group))
        if key not in cls._cache:
            # every rank sees the same sequence of requests, so every rank evicts the same entry
            while len(cls._cache) >= cls._cache_max:
                cls._cache.pop(next(iter(cls._cache)))
            cls._cache[key] = cls(num_edges, device, group)
        return cls._cache[key]


def gauss_newton_sharded(prm, Twc, Xs, Cs, ii, jj, idx_local, valid_local, Q_local, edge_begin,
                         max_iter, delta_thresh, group=None, edge_counts=None, p2p=False, check=True,
                         packed_tail=None):
    """Edge-sharded solve: each rank accumulates its contiguous slice of the directed edges
    (`shard_edges`), the 56-double per-edge blocks are exchanged — with ONE all-gather per iteration
    (`exchange_blocks`) or, p2p=True, inside the kernels over NVLink peer memory — then every rank
    runs the identical replicated assembly + Cholesky + retraction, so poses stay bit-identical on
    all ranks — and identical to the single-GPU solve, because an edge's block does not depend on
    which rank computed it.  The local slice is in the reference layout, or packed records when
    valid_local and Q_local are None (see GNSolver).
    check=True (default): after the loop the status words are read back (one 16-byte synchronising
    read) and a peer-memory timeout raises RuntimeError on EVERY rank — a rank that stopped hearing
    from a peer stops signalling, so its peers time out as well; nobody keeps diverged poses silently."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    peers = PeerBlocks.get(ii.shape[0], Twc.device, group) if (p2p and world > 1) else None
    solver = GNSolver(prm, Twc, Xs, Cs, ii, jj, idx_local, valid_local, Q_local, edge_begin, peers=peers,
                      packed_tail=packed_tail)
    if world > 1 and edge_counts is None and peers is None:
        raise RuntimeError("gauss_newton_sharded: edge_counts (from shard_edges) is required")
    for _ in range(int(max_iter)):
        solver.accumulate()
        if world > 1 and peers is None:
            exchange_blocks(solver.blocks, solver.edge_begin, solver.Ml, list(edge_counts), group)
        solver.solve_step(delta_thresh)
    if check and peers is not None and int(max_iter) > 0:
        st = solver.status()
        if st["peer_timeout"]:
            raise RuntimeError("gauss_newton_sharded: a peer's blocks did not arrive within the device-side "
                               f"timeout after {st['iterations']} iteration(s); the poses of this solve are not "
                               "valid (all ranks raise)")
    return [solver.dx]


# --------------------------------------------------------------------------- tracker

def track_ray_dist(Xf, Xk, Qk, valid, T_CkCf, sigma_ray, sigma_dist, huber_k, max_iters, rel_error,
                   delta_norm, want_cost=False):
    """Whole opt_pose_ray_dist_sim3 loop (tracker.py:173-214) on the relative pose.
    T_CkCf: (8,) f32 CUDA tensor updated in place.  -> status (4,) i32 CUDA tensor
    [iterations, converged, cholesky_failed, 0] (and the per-iteration cost if requested)."""
    Xf = _req(Xf, "Xf", torch.float32, 2)
    Xk = _req(Xk, "Xk", torch.float32, 2)
    Qk = _req(Qk, "Qk", torch.float32)
    valid = _req(valid, "valid", torch.bool)
    T = _req(T_CkCf, "T_CkCf", torch.float32)
    n = Xf.shape[0]
    with torch.cuda.device(Xf.device):
        status = torch.zeros((4,), dtype=torch.int32, device=Xf.device)
        cost = torch.zeros((max(int(max_iters), 1),), dtype=torch.float32, device=Xf.device) if want_cost else None
        nbytes = _lib.m3r_track_workspace_bytes(n)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=Xf.device)
        _check(_lib.m3r_track_ray_dist(_ptr(Xf), _ptr(Xk), _ptr(Qk), _ptr(valid), n, float(sigma_ray),
                                       float(sigma_dist), float(huber_k), int(max_iters),
                                       float(rel_error), float(delta_norm), _ptr(T), _ptr(status),
                                       _ptr(cost), _ptr(ws), nbytes, _stream(Xf)), "track_ray_dist")
    return (status, cost) if want_cost else status


TRACK_NOALLOC_MAX_N = 1 << 22
ETOOLARGE = -4


def track_ray_dist_noalloc(Xf, Xk, Qk, valid, T_CkCf, status, sigma_ray, sigma_dist, huber_k, max_iters,
                           rel_error, delta_norm, cost=None, n=None):
    """m3r_track_ray_dist_noalloc: track_ray_dist for callers that cannot allocate.  The caller
    passes every output (`status` (4,) i32, optionally `cost` (max_iters,) f32); the scratch is a
    static device array inside the library, so neither this wrapper nor the call allocates device
    memory.  No counterpart in the reference.
    -> the C return code (0 ok, ETOOLARGE beyond TRACK_NOALLOC_MAX_N points); other errors raise.
    Calls that overlap on one GPU share the scratch and race."""
    n = int(Xf.shape[0]) if n is None else int(n)
    rc = _lib.m3r_track_ray_dist_noalloc(_ptr(Xf), _ptr(Xk), _ptr(Qk), _ptr(valid), n, float(sigma_ray),
                                         float(sigma_dist), float(huber_k), int(max_iters), float(rel_error),
                                         float(delta_norm), _ptr(T_CkCf), _ptr(status), _ptr(cost), _stream(Xf))
    if rc not in (0, ETOOLARGE):
        _check(rc, "track_ray_dist_noalloc")
    return rc


def track_calib(Xf, meas_k, Qk, valid, valid_meas_k, K, img_size, T_CkCf, sigma_pixel, sigma_depth,
                pixel_border, z_eps, huber_k, max_iters, rel_error, delta_norm, want_cost=False):
    """Whole opt_pose_calib_sim3 loop (tracker.py:216-266) on the relative pose; arguments as
    track_ray_dist plus meas_k (n,3) f32, valid_meas_k (n,1) bool, K (3,3), img_size (h, w)."""
    Xf = _req(Xf, "Xf", torch.float32, 2)
    meas_k = _req(meas_k, "meas_k", torch.float32, 2)
    Qk = _req(Qk, "Qk", torch.float32)
    valid = _req(valid, "valid", torch.bool)
    valid_meas_k = _req(valid_meas_k, "valid_meas_k", torch.bool)
    T = _req(T_CkCf, "T_CkCf", torch.float32)
    Kd = K.detach().to(device=Xf.device, dtype=torch.float32).contiguous()  # no-op when already there
    if Kd.shape != (3, 3):
        raise RuntimeError("K must be (3,3)")
    n = Xf.shape[0]
    with torch.cuda.device(Xf.device):
        status = torch.zeros((4,), dtype=torch.int32, device=Xf.device)
        cost = torch.zeros((max(int(max_iters), 1),), dtype=torch.float32, device=Xf.device) if want_cost else None
        nbytes = _lib.m3r_track_workspace_bytes(n)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=Xf.device)
        _check(_lib.m3r_track_calib(_ptr(Xf), _ptr(meas_k), _ptr(Qk), _ptr(valid), _ptr(valid_meas_k), n,
                                    _ptr(Kd), int(img_size[0]), int(img_size[1]), int(pixel_border), float(z_eps),
                                    float(sigma_pixel), float(sigma_depth), float(huber_k), int(max_iters),
                                    float(rel_error), float(delta_norm), _ptr(T), _ptr(status), _ptr(cost),
                                    _ptr(ws), nbytes, _stream(Xf)), "track_calib")
    return (status, cost) if want_cost else status


# --------------------------------------------------------------------------- front-end steps

def edge_build(idx_src2dst, valid_match, Q_src, Q_dst, Q_conf):
    """One direction of add_factors' edge tensors (global_opt.py:55-67).
    -> (Q (E,n,1) f32 = sqrt(Q_src[idx] * Q_dst), valid_count (E,) i32 of valid_match & (Q > Q_conf))."""
    idx = _req(idx_src2dst, "idx", torch.int64, 2)
    vm = _req(valid_match, "valid_match", torch.bool)
    Qs = _req(Q_src, "Q_src", torch.float32)
    Qd = _req(Q_dst, "Q_dst", torch.float32)
    E, n = idx.shape
    with torch.cuda.device(idx.device):
        Q = torch.empty((E, n, 1), dtype=torch.float32, device=idx.device)
        cnt = torch.empty((E,), dtype=torch.int32, device=idx.device)
        _check(_lib.m3r_edge_build(_ptr(idx), _ptr(vm), _ptr(Qs), _ptr(Qd), E, n, float(Q_conf), _ptr(Q),
                                   _ptr(cnt), _stream(idx)), "edge_build")
    return Q, cnt


def edge_build_packed(idx_src2dst, valid_match, Q_src, Q_dst, Q_conf, out=None, want_Q=False):
    """edge_build writing PACKED records (m3r_edge_build_packed): -> (pk (E,n) int64, valid_count (E,) i32
    [, Q (E,n,1) f32]).  `out`: an (E,n) int64 destination (e.g. the tail rows of FactorGraph's store)."""
    idx = _req(idx_src2dst, "idx", torch.int64, 2)
    vm = _req(valid_match, "valid_match", torch.bool)
    Qs = _req(Q_src, "Q_src", torch.float32)
    Qd = _req(Q_dst, "Q_dst", torch.float32)
    E, n = idx.shape
    with torch.cuda.device(idx.device):
        if out is None:
            out = torch.empty((E, n), dtype=torch.int64, device=idx.device)
        else:
            out = _req_packed(out, "out", n)
            if out.shape[0] != E:
                raise RuntimeError("edge_build_packed: out has the wrong number of rows")
        Q = torch.empty((E, n, 1), dtype=torch.float32, device=idx.device) if want_Q else None
        cnt = torch.empty((E,), dtype=torch.int32, device=idx.device)
        _check(_lib.m3r_edge_build_packed(_ptr(idx), _ptr(vm), _ptr(Qs), _ptr(Qd), E, n, float(Q_conf), _ptr(out),
                                          _ptr(Q), _ptr(cnt), _stream(idx)), "edge_build_packed")
    return (out, cnt, Q) if want_Q else (out, cnt)


POINTMAP_MODES = {"weighted_pointmap": 0, "weighted_spherical": 1, "indep_conf": 2}


def pointmap_update(X_canon, C, X_new, C_new, mode="weighted_pointmap", T_act=None):
    """In place: Frame.update_pointmap's per-point modes (frame.py:69-101); T_act (8,) f32 applies
    Sim3.act to X_new first (tracker.py:97-98)."""
    Xc = _req(X_canon, "X_canon", torch.float32, 2)
    Cc = _req(C, "C", torch.float32)
    Xn = _req(X_new, "X_new", torch.float32, 2)
    Cn = _req(C_new, "C_new", torch.float32)
    if T_act is not None:
        T_act = _req(T_act, "T_act", torch.float32)
    if Xn.shape != Xc.shape or Cn.numel() != Cc.numel() or Cc.numel() != Xc.shape[0]:
        raise RuntimeError("pointmap_update: shape mismatch")
    with torch.cuda.device(Xc.device):
        _check(_lib.m3r_pointmap_update(_ptr(Xc), _ptr(Cc), _ptr(Xn), _ptr(Cn), _ptr(T_act), Xc.shape[0],
                                        POINTMAP_MODES[mode], _stream(Xc)), "pointmap_update")
    return Xc, Cc


# --------------------------------------------------------------------------- headless viewer splat

def render_opts(**kw):
    """m3r_render_opts with the shaders' uniform defaults (surfelmap.glsl:17,99-104,
    trianglemap.glsl:17-18) and the viewer's state (visualization.py:38,77), overridden by kw."""
    o = _RenderOpts()
    _lib.m3r_render_default_opts(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(f"render_opts: unknown option {k}")
        if k in ("lightpos", "phong", "base_color"):
            getattr(o, k)[:] = [float(x) for x in v]
        else:
            setattr(o, k, v)
    return o


def render_clear(fb, clear_rgba=0xFF26231E):
    """fb (height,width) int64 CUDA (packed depth|RGBA8 words): depth 1.0 + clear colour
    (default #1E2326, visualization.py:53)."""
    fb = _req(fb, "fb", torch.int64, 2)
    _check(_lib.m3r_render_clear(_ptr(fb), fb.shape[1], fb.shape[0], C.c_uint32(clear_rgba), _stream(fb)),
           "render_clear")
    return fb


def render_pointmaps(program, X, Cf, img, models, camera, proj, fb, opts=None):
    """Splat K keyframe pointmaps into fb with the arithmetic of `program` ("surfelmap" |
    "trianglemap": resources/programs/*.glsl).  X (K,H,W,3), Cf (K,H,W), img (K,H,W,3), models
    (K,4,4) f32 CUDA; camera, proj: 16 floats each, row-major, host."""
    X = _req(X, "X", torch.float32, 4)
    Cf = _req(Cf, "C", torch.float32, 3)
    img = _req(img, "img", torch.float32, 4)
    models = _req(models, "models", torch.float32, 3)
    fb = _req(fb, "fb", torch.int64, 2)
    K, H, W, _ = X.shape
    if Cf.shape != (K, H, W) or img.shape != (K, H, W, 3) or models.shape != (K, 4, 4) or X.shape[3] != 3:
        raise RuntimeError("render_pointmaps: X (K,H,W,3), C (K,H,W), img (K,H,W,3), models (K,4,4)")
    if program not in ("surfelmap", "trianglemap"):
        raise RuntimeError("render_pointmaps: program must be 'surfelmap' or 'trianglemap'")
    view = _RenderView()
    cam = [float(x) for x in torch.as_tensor(camera, dtype=torch.float32).reshape(-1).tolist()]
    prj = [float(x) for x in torch.as_tensor(proj, dtype=torch.float32).reshape(-1).tolist()]
    if len(cam) != 16 or len(prj) != 16:
        raise RuntimeError("render_pointmaps: camera and proj must be 4x4")
    view.camera[:] = cam
    view.proj[:] = prj
    view.width, view.height = fb.shape[1], fb.shape[0]
    o = opts if opts is not None else render_opts()
    fn = _lib.m3r_render_surfels if program == "surfelmap" else _lib.m3r_render_triangles
    with torch.cuda.device(X.device):
        _check(fn(_ptr(X), _ptr(Cf), _ptr(img), _ptr(models), K, H, W, C.byref(view), C.byref(o),
                  _ptr(fb), _stream(X)), "render_pointmaps")
    return fb


def render_resolve(fb, want_depth=True):
    """-> (rgba (h,w,4) u8 top-down, depth (h,w) f32 top-down or None)"""
    fb = _req(fb, "fb", torch.int64, 2)
    h, w = fb.shape
    with torch.cuda.device(fb.device):
        rgba = torch.empty((h, w, 4), dtype=torch.uint8, device=fb.device)
        depth = torch.empty((h, w), dtype=torch.float32, device=fb.device) if want_depth else None
        _check(_lib.m3r_render_resolve(_ptr(fb), w, h, _ptr(rgba), _ptr(depth), _stream(fb)), "render_resolve")
    return rgba, depth
