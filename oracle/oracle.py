"""numpy front-end of oracle/libm3r_oracle.so — the CPU restatement of the reference hot path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
Argument names and order follow the reference Python API (gn.cpp:116-122).
"""
import ctypes as C
import os

import numpy as np

from . import build_oracle

_lib = None


class GNParams(C.Structure):
    _fields_ = [("mode", C.c_int), ("sigma_a", C.c_float), ("sigma_b", C.c_float),
                ("C_thresh", C.c_float), ("Q_thresh", C.c_float),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("height", C.c_int), ("width", C.c_int), ("pixel_border", C.c_int),
                ("z_eps", C.c_float)]


MODE_RAYS, MODE_CALIB, MODE_POINTS = 0, 1, 2


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle.build())
        _lib.orc_h2f.restype = C.c_float
        _lib.orc_h2f.argtypes = [C.c_uint16]
        _lib.orc_f2h.restype = C.c_uint16
        _lib.orc_f2h.argtypes = [C.c_float]
        _lib.orc_gauss_newton.restype = C.c_int
        _lib.orc_sparse_block_solve.restype = C.c_int
    return _lib


def set_threads(k):
    """OpenMP thread count for the oracle (cpu_baseline reports it as `cores`)."""
    os.environ["OMP_NUM_THREADS"] = str(k)
    try:
        C.CDLL("libgomp.so.1").omp_set_num_threads(int(k))
    except OSError:
        pass


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def f2h(x):
    x = _c(x, np.float32)
    out = np.empty(x.shape, np.uint16)
    lib().orc_f2h_array(_p(x), _p(out), C.c_int64(x.size))
    return out


def h2f(x):
    x = _c(x, np.uint16)
    out = np.empty(x.shape, np.float32)
    lib().orc_h2f_array(_p(x), _p(out), C.c_int64(x.size))
    return out


def refine_matches(D11, D21, p1, radius, dilation_max):
    """D11 (b,H,W,F), D21 (b,n,F): float16 (or uint16 bit patterns) or float32; p1 (b,n,2) int64."""
    p1 = _c(p1, np.int64)
    b, H, W, F = D11.shape
    n = p1.shape[1]
    out = np.zeros_like(p1)
    if D11.dtype == np.float32:
        D11 = _c(D11, np.float32)
        D21 = _c(D21, np.float32)
        fn = lib().orc_refine_matches_f32
    else:
        D11 = np.ascontiguousarray(D11).view(np.uint16)
        D21 = np.ascontiguousarray(D21).view(np.uint16)
        fn = lib().orc_refine_matches_f16
    fn(_p(D11), _p(D21), _p(p1), b, H, W, n, F, int(radius), int(dilation_max), _p(out))
    return out


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    img = _c(rays_img_with_grad, np.float32)
    pts = _c(pts_3d_norm, np.float32)
    p0 = _c(p_init, np.float32)
    b, H, W, c = img.shape
    assert c == 9
    n = p0.shape[1]
    p_new = np.zeros((b, n, 2), np.float32)
    conv = np.zeros((b, n), np.uint8)
    lib().orc_iter_proj(_p(img), _p(pts), _p(p0), b, H, W, n, int(max_iter),
                        C.c_float(lambda_init), C.c_float(cost_thresh), _p(p_new), _p(conv))
    return p_new, conv.astype(bool)


def pose_retr(poses, dx, num_fix=1):
    poses = _c(poses, np.float32).copy()
    dx = _c(dx, np.float32)
    lib().orc_pose_retr(_p(poses), _p(dx), poses.shape[0], int(num_fix))
    return poses


def rel_sim3(Ti, Tj):
    Ti, Tj = _c(Ti, np.float32), _c(Tj, np.float32)
    out = np.zeros(8, np.float32)
    lib().orc_relSim3(_p(Ti), _p(Tj), _p(out))
    return out


def act_sim3(T, X):
    T, X = _c(T, np.float32), _c(X, np.float32)
    out = np.zeros(3, np.float32)
    lib().orc_actSim3(_p(T), _p(X), _p(out))
    return out


def adj_inv_row(T, X):
    T, X = _c(T, np.float32), _c(X, np.float32)
    out = np.zeros(7, np.float32)
    lib().orc_adj_inv_row(_p(T), _p(X), _p(out))
    return out


def make_params(mode, sigma_a, sigma_b=0.0, C_thresh=0.0, Q_thresh=1.5, K=None, height=0,
                width=0, pixel_border=0, z_eps=0.0):
    p = GNParams()
    p.mode = mode
    p.sigma_a, p.sigma_b, p.C_thresh, p.Q_thresh = sigma_a, sigma_b, C_thresh, Q_thresh
    if K is not None:
        K = np.asarray(K, np.float32)
        p.fx, p.fy, p.cx, p.cy = float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])
    p.height, p.width, p.pixel_border, p.z_eps = int(height), int(width), int(pixel_border), z_eps
    return p


def edge_rows(ii, jj, num_fix=1):
    """get_unique_kf_idx + create_inds (gn_kernels.cu:161-170)."""
    ii = np.asarray(ii, np.int64)
    jj = np.asarray(jj, np.int64)
    uniq = np.unique(np.concatenate([ii, jj]))
    ie = np.searchsorted(uniq, ii).astype(np.int64)
    je = np.searchsorted(uniq, jj).astype(np.int64)
    return uniq, ie, je, ie - num_fix, je - num_fix


def gn_blocks(prm, Twc, Xs, Cs, ii_edge, jj_edge, idx_ii2jj, valid_match, Q, want64=False):
    """One accumulation pass -> Hs (4,M,7,7), gs (2,M,7) in the reference layout."""
    Twc, Xs, Cs = _c(Twc, np.float32), _c(Xs, np.float32), _c(Cs, np.float32)
    ie, je = _c(ii_edge, np.int64), _c(jj_edge, np.int64)
    idx = _c(idx_ii2jj, np.int64)
    vm = _c(valid_match, np.uint8)
    Q = _c(Q, np.float32)
    M, n = idx.shape[0], Xs.shape[1]
    Hs = np.zeros((4, M, 7, 7), np.float32)
    gs = np.zeros((2, M, 7), np.float32)
    Hs64 = np.zeros((4, M, 7, 7), np.float64) if want64 else None
    gs64 = np.zeros((2, M, 7), np.float64) if want64 else None
    lib().orc_gn_blocks(C.byref(prm), _p(Twc), _p(Xs), _p(Cs), _p(ie), _p(je), _p(idx), _p(vm),
                        _p(Q), n, M, _p(Hs), _p(gs),
                        _p(Hs64) if want64 else None, _p(gs64) if want64 else None)
    return (Hs, gs, Hs64, gs64) if want64 else (Hs, gs)


def sparse_block_solve(Hs, gs, ii_opt, jj_opt, num_free, want_system=False):
    Hs, gs = _c(Hs, np.float32), _c(gs, np.float32)
    io, jo = _c(ii_opt, np.int64), _c(jj_opt, np.int64)
    M = io.shape[0]
    D = 7 * num_free
    dx = np.zeros((num_free, 7), np.float32)
    A = np.zeros((D, D), np.float64) if want_system else None
    bv = np.zeros((D,), np.float64) if want_system else None
    ok = lib().orc_sparse_block_solve(_p(Hs), _p(gs), _p(io), _p(jo), M, num_free,
                                      _p(A) if want_system else None,
                                      _p(bv) if want_system else None, _p(dx))
    return (dx, bool(ok), A, bv) if want_system else (dx, bool(ok))


def gauss_newton(prm, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh,
                 history=False):
    """Full loop.  Returns (Twc_new, dx, iterations[, dx_hist]); the input Twc is not mutated."""
    Twc = _c(Twc, np.float32).copy()
    Xs, Cs = _c(Xs, np.float32), _c(Cs, np.float32)
    ii, jj = _c(ii, np.int64), _c(jj, np.int64)
    idx = _c(idx_ii2jj, np.int64)
    vm = _c(valid_match, np.uint8)
    Q = _c(Q, np.float32)
    N, n = Xs.shape[0], Xs.shape[1]
    M = ii.shape[0]
    dx = np.zeros((N - 1, 7), np.float32)
    hist = np.zeros((max(max_iter, 1), N - 1, 7), np.float32) if history else None
    its = lib().orc_gauss_newton(C.byref(prm), _p(Twc), _p(Xs), _p(Cs), _p(ii), _p(jj), _p(idx),
                                 _p(vm), _p(Q), N, n, M, int(max_iter), C.c_float(delta_thresh),
                                 _p(dx), _p(hist) if history else None)
    return (Twc, dx, its, hist[:its]) if history else (Twc, dx, its)
