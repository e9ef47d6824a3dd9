"""Guard-band runs of every C-ABI entry point (compute-sanitizer is closed on this pool, so the
bounds are checked by hand).

Each operation runs twice on the same seeded inputs, at sizes that are multiples of nothing
(37x53 images, 1001 queries): once on ordinary torch allocations, once with
  * every INPUT copied into the middle of its own buffer whose 4 KB margins hold NaN (float
    tensors) or 0 (index / flag tensors) — a read past either end that reaches the result makes
    the two runs differ;
  * every OUTPUT and WORKSPACE the host mirror allocates carved out of a buffer whose margins
    hold a canary byte and whose body is pre-filled with 0xFF — a write outside the tensor breaks
    the canary, an element the kernel never writes (or a workspace word it reads before writing)
    shows up as NaN / -1 in the comparison.
The two runs must agree bit for bit and every margin must be untouched afterwards.
"""
import math

import pytest
import torch

from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PAD = 4096          # bytes on either side; keeps the 256-B alignment torch allocations have
CANARY = 0xA5
H, W = 37, 53       # odd on purpose: no tile, vector or warp size divides them


class Arena:
    def __init__(self):
        self.items = []  # (raw buffer, body bytes, margin byte)

    def _carve(self, shape, dtype, device, margin, body):
        shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)))
        nbytes = math.prod(shape) * torch.empty((), dtype=dtype).element_size()
        raw = torch.full((PAD + nbytes + PAD,), margin, dtype=torch.uint8, device=device)
        raw[PAD:PAD + nbytes] = body
        self.items.append((raw, nbytes, margin))
        return raw[PAD:PAD + nbytes].view(dtype).view(shape)

    def out(self, shape, dtype, device, body=0xFF):
        return self._carve(shape, dtype, device, CANARY, body)

    def inp(self, t):
        """Copy of input `t` with hostile margins (NaN for floats, 0 for integers and flags)."""
        if t is None or not isinstance(t, torch.Tensor) or not t.is_cuda:
            return t
        g = self._carve(t.shape, t.dtype, t.device, 0xFF if t.dtype.is_floating_point else 0x00, 0)
        g.copy_(t)
        return g

    def check(self):
        torch.cuda.synchronize()
        for k, (raw, nbytes, margin) in enumerate(self.items):
            lo, hi = raw[:PAD], raw[PAD + nbytes:]
            assert bool((lo == margin).all()), f"buffer {k} ({nbytes} B): write below its start"
            assert bool((hi == margin).all()), f"buffer {k} ({nbytes} B): write past its end"


class GuardedTorch:
    """Stands in for the `torch` name inside mast3r_slam_backends: allocations come from the arena."""

    def __init__(self, arena):
        self._arena = arena

    def __getattr__(self, name):
        return getattr(torch, name)

    def empty(self, shape, dtype=torch.float32, device=None):
        return self._arena.out(shape, dtype, device)

    def zeros(self, shape, dtype=torch.float32, device=None):
        return self._arena.out(shape, dtype, device, body=0)

    def empty_like(self, t):
        return self._arena.out(t.shape, t.dtype, t.device)


def _bits(t):
    return t.contiguous().view(-1).view(torch.uint8)


def _same(a, b, what):
    if a is None and b is None:
        return
    assert a.shape == b.shape and a.dtype == b.dtype, what
    assert torch.equal(_bits(a), _bits(b)), f"{what}: guarded run differs from the plain run"


def _both(backends, monkeypatch, fn, inputs):
    """fn(*tensors) -> tuple of tensors (outputs and tensors updated in place)."""
    plain = fn(*[t.clone() if isinstance(t, torch.Tensor) else t for t in inputs])
    torch.cuda.synchronize()
    arena = Arena()
    guarded_in = [arena.inp(t) for t in inputs]
    with monkeypatch.context() as m:
        m.setattr(backends, "torch", GuardedTorch(arena))
        guarded = fn(*guarded_in)
    arena.check()
    assert len(plain) == len(guarded)
    for k, (a, b) in enumerate(zip(plain, guarded)):
        _same(a, b, f"{getattr(fn, '__name__', 'op')} result {k}")


def test_matching_entry_points(backends, monkeypatch):
    b, F = 2, 24
    m = {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v)
         for k, v in synth.make_match_inputs(b, H, W, F, seed=91, device="cpu").items()}
    X11, X21, D11, D21 = m["X11"], m["X21"], m["D11"], m["D21"]
    n = H * W

    for packed in (False, True):
        _both(backends, monkeypatch, lambda x, d, p=packed: backends.match_prep(x, d, packed=p), [X11, D11])
    _both(backends, monkeypatch, lambda x: backends.img_gradient(x), [X11.permute(0, 3, 1, 2).contiguous()])
    rays9, D11h = backends.match_prep(X11, D11)
    rays12, _ = backends.match_prep(X11, None, packed=True)
    pts = torch.nn.functional.normalize(X21.view(b, n, 3), dim=-1).contiguous()
    p_init = torch.stack(torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")[::-1],
                         -1).view(1, n, 2).expand(b, n, 2).float().contiguous()
    _both(backends, monkeypatch, lambda r, p, q: tuple(backends.iter_proj(r, p, q, 10, 1e-8, 1e-6)),
          [rays9, pts, p_init])
    idx_init = torch.randint(0, n, (b, n), device=DEV, generator=torch.Generator(DEV).manual_seed(5))
    for rays in (rays9, rays12):
        for init in (None, idx_init):
            _both(backends, monkeypatch,
                  lambda r, a, c, i: backends.match_iter_proj_fused(r, a, c, i, 10, 1e-8, 1e-6, 0.1, want_p_new=True),
                  [rays, X11, X21, init])
    p1, _ = backends.match_iter_proj_fused(rays12, X11, X21, None, 10, 1e-8, 1e-6, 0.1)
    _both(backends, monkeypatch, lambda d, q, p: (backends.match_refine_fused(d, q, p, 3, 5),),
          [D11h, D21.reshape(b, n, F).contiguous(), p1])
    # the reference entry point, ragged query count, both descriptor types, window clipped at every border
    nq = 1001
    q1 = torch.stack([torch.randint(0, W, (b, nq), device=DEV), torch.randint(0, H, (b, nq), device=DEV)], -1)
    q1[:, :4] = torch.tensor([[0, 0], [W - 1, 0], [0, H - 1], [W - 1, H - 1]], device=DEV)
    D21q = D21.reshape(b, n, F)[:, :nq].contiguous()
    for cast in (torch.float16, torch.float32):
        for (radius, dil) in ((3, 5), (2, 1), (0, 1)):
            _both(backends, monkeypatch, lambda d, q, p, r=radius, s=dil: tuple(backends.refine_matches(d, q, p, r, s)),
                  [D11.to(cast), D21q.to(cast), q1])


def _graph(calib=False):
    g = synth.make_graph(N=5, counts={1: None, 2: 2, "random": 1}, H=H, W=W, seed=1290, device=DEV)
    if calib:
        g["Xs"] = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])
    return g


@pytest.mark.parametrize("mode", ["rays", "calib", "points"])
def test_gauss_newton_entry_points(backends, monkeypatch, mode):
    g = _graph(calib=(mode == "calib"))

    def run(Twc, Xs, Cs, ii, jj, idx, valid, Q):
        if mode == "rays":
            (dx,) = backends.gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, 0.003, 10.0, 0.0, 1.5, 3, 0.0)
        elif mode == "calib":
            (dx,) = backends.gauss_newton_calib(Twc, Xs, Cs, g["K"], ii, jj, idx, valid, Q, H, W, -10, 1e-6,
                                                1.0, 10.0, 0.0, 1.5, 3, 0.0)
        else:
            (dx,) = backends.gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, 0.05, 0.0, 1.5, 3, 0.0)
        return dx, Twc

    _both(backends, monkeypatch, run, [g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"]])


def test_stepwise_solver_and_pose_retr(backends, monkeypatch):
    g = _graph()
    prm = backends._params(backends.GN_RAYS, 0.003, 10.0, 0.0, 1.5)

    def run(Twc, Xs, Cs, ii, jj, idx, valid, Q):
        s = backends.GNSolver(prm, Twc, Xs, Cs, ii, jj, idx, valid, Q)
        blocks = s.accumulate().clone()
        s.solve_step(0.0)
        return blocks, s.dx, Twc

    _both(backends, monkeypatch, run, [g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"]])
    dx = 0.05 * torch.randn((4, 7), device=DEV, generator=torch.Generator(DEV).manual_seed(3))
    _both(backends, monkeypatch, lambda P, d: (backends.pose_retr(P, d),), [g["Twc"], dx])


def test_tracker_entry_points(backends, monkeypatch):
    d = synth.make_tracker_inputs(H, W, seed=7, device=DEV)
    T = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float32, device=DEV)

    def ray(Xf, Xk, Qk, valid, T):
        status, cost = backends.track_ray_dist(Xf, Xk, Qk, valid, T, 0.003, 10.0, 1.345, 50, 1e-3, 1e-3,
                                               want_cost=True)
        return status, cost, T

    _both(backends, monkeypatch, ray, [d["Xf"], d["Xk"], d["Qk"], d["valid"], T])
    c = synth.make_tracker_calib_inputs(H, W, seed=8, device=DEV)

    def calib(Xf, meas, Qk, valid, vm, T):
        status, cost = backends.track_calib(Xf, meas, Qk, valid, vm, c["K"], (H, W), T, 1.0, 10.0, -10, 1e-6,
                                            1.345, 50, 1e-3, 1e-3, want_cost=True)
        return status, cost, T

    _both(backends, monkeypatch, calib, [c["Xf"], c["meas_k"], c["Qk"], c["valid"], c["valid_meas_k"], T])


def test_frontend_entry_points(backends, monkeypatch):
    for n in (1001, 1000):  # scalar kernel / 4-points-per-thread kernel
        gen = torch.Generator(DEV).manual_seed(n)
        E = 3
        idx = torch.randint(0, n, (E, n), device=DEV, generator=gen)
        vm = torch.rand((E, n, 1), device=DEV, generator=gen) > 0.3
        Qs = 1.0 + torch.rand((E, n, 1), device=DEV, generator=gen)
        Qd = 1.0 + torch.rand((E, n, 1), device=DEV, generator=gen)
        _both(backends, monkeypatch, lambda i, v, a, b: backends.edge_build(i, v, a, b, 1.5), [idx, vm, Qs, Qd])
        Xc = torch.randn((n, 3), device=DEV, generator=gen)
        Xn = torch.randn((n, 3), device=DEV, generator=gen)
        Cc = 1.0 + torch.rand((n, 1), device=DEV, generator=gen)
        Cn = 1.0 + torch.rand((n, 1), device=DEV, generator=gen)
        T = torch.tensor([0.1, -0.2, 0.05, 0.0, 0.0, math.sin(0.1), math.cos(0.1), 1.1], device=DEV)
        for mode in backends.POINTMAP_MODES:
            for T_act in (None, T):
                _both(backends, monkeypatch,
                      lambda a, b, c, d, t, md=mode: backends.pointmap_update(a, b, c, d, md, t),
                      [Xc, Cc, Xn, Cn, T_act])


@pytest.mark.parametrize("program", ["surfelmap", "trianglemap"])
def test_render_entry_points(backends, monkeypatch, program):
    sc = synth.make_render_scene(K=2, H=H, W=W, seed=1243)
    w, h = 101, 67

    def run(X, Cf, img, models, fb):
        backends.render_clear(fb)
        backends.render_pointmaps(program, X, Cf, img, models, sc["camera"], sc["proj"], fb,
                                  backends.render_opts(radius=0.02))
        rgba, depth = backends.render_resolve(fb)
        return fb, rgba, depth

    fb = torch.zeros((h, w), dtype=torch.int64, device=DEV)
    _both(backends, monkeypatch, run, [sc["X"].to(DEV), sc["C"].to(DEV), sc["img"].to(DEV), sc["models"].to(DEV), fb])


def test_guard_detects_planted_violations():
    """The checker itself: a one-byte overrun, an underrun and an unwritten output are all seen."""
    for where in (-1, None):
        arena = Arena()
        t = arena.out((10,), torch.float32, DEV)
        raw = arena.items[0][0]
        raw[PAD - 1 if where == -1 else PAD + 40] = 0
        with pytest.raises(AssertionError):
            arena.check()
    arena = Arena()
    t = arena.out((4,), torch.float32, DEV)
    t[:3] = 1.0  # "kernel" forgets the last element
    arena.check()
    with pytest.raises(AssertionError):
        _same(torch.ones(4, device=DEV), t, "planted")
