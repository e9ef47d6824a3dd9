"""Headless viewer splat (csrc/m3r_render.cu) against the CPU restatement of the shaders
(oracle/m3r_render_oracle.c).  Every operation in both is a single IEEE rounding in the same
order, and the depth test is an order-independent 64-bit min, so the framebuffers must be
bit-identical.  (Parity with OpenGL itself is unpinned: no graphics API on this box.)"""
import numpy as np
import pytest
import torch

from oracle import render as R
from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _gpu_render(be, program, sc, w, h, **kw):
    fb = torch.empty((h, w), dtype=torch.int64, device=DEV)
    be.render_clear(fb)
    be.render_pointmaps(program, sc["X"].to(DEV), sc["C"].to(DEV), sc["img"].to(DEV), sc["models"].to(DEV),
                        sc["camera"], sc["proj"], fb, be.render_opts(**kw))
    torch.cuda.synchronize()
    return fb


def _cpu_render(program, sc, w, h, **kw):
    return R.render(program, sc["X"].numpy(), sc["C"].numpy(), sc["img"].numpy(), sc["models"].numpy(),
                    sc["camera"].numpy(), sc["proj"].numpy(), w, h, **kw)


@pytest.mark.parametrize("program,kw", [
    ("surfelmap", dict(radius=0.008)),
    ("surfelmap", dict(radius=0.012, use_img=0)),
    ("surfelmap", dict(radius=0.008, show_normal=1, cull_face=0)),
    ("trianglemap", dict()),
    ("trianglemap", dict(depth_bias=0.001, use_img=0)),
    ("trianglemap", dict(slant_threshold=0.3, show_normal=1)),
])
def test_splat_bit_identical_to_oracle(backends, program, kw):
    sc = synth.make_render_scene(K=3, H=96, W=128, seed=1240)
    w, h = 320, 240
    fb = _gpu_render(backends, program, sc, w, h, **kw).cpu().numpy().view(np.uint64)
    ref = _cpu_render(program, sc, w, h, **kw)
    covered = (ref >> np.uint64(32)) != np.uint64(0x3F800000)
    assert covered.mean() > 0.2, "scene does not fill the view"
    assert np.array_equal(fb, ref), f"{(fb != ref).sum()} of {fb.size} framebuffer words differ"


def test_splat_edge_cases(backends):
    sc = synth.make_render_scene(K=2, H=48, W=64, seed=1241)
    # zero keyframes: nothing drawn, no launch error
    fb = torch.empty((60, 80), dtype=torch.int64, device=DEV)
    backends.render_clear(fb, 0xFF000000)
    empty = {k: (v[:0] if isinstance(v, torch.Tensor) and k in ("X", "C", "img", "models") else v) for k, v in sc.items()}
    backends.render_pointmaps("surfelmap", empty["X"].to(DEV), empty["C"].to(DEV), empty["img"].to(DEV),
                              empty["models"].to(DEV), sc["camera"], sc["proj"], fb)
    torch.cuda.synchronize()
    assert (fb.cpu().numpy().view(np.uint64) == np.uint64((0x3F800000 << 32) | 0xFF000000)).all()
    # a camera that looks away: every vertex has w <= 0 -> dropped
    cam_away = torch.diag(torch.tensor([1.0, 1.0, -1.0, 1.0])) @ sc["camera"]
    sc2 = dict(sc, camera=cam_away.contiguous())
    a = _gpu_render(backends, "trianglemap", sc2, 80, 60).cpu().numpy().view(np.uint64)
    assert np.array_equal(a, _cpu_render("trianglemap", sc2, 80, 60))
    # target much larger than the pointmap footprint / 1-pixel target / non-multiple sizes
    for (w, h) in [(1, 1), (37, 23), (401, 299)]:
        g = _gpu_render(backends, "trianglemap", sc, w, h).cpu().numpy().view(np.uint64)
        assert np.array_equal(g, _cpu_render("trianglemap", sc, w, h)), (w, h)
    # usage errors
    with pytest.raises(RuntimeError):
        backends.render_pointmaps("nope", sc["X"].to(DEV), sc["C"].to(DEV), sc["img"].to(DEV),
                                  sc["models"].to(DEV), sc["camera"], sc["proj"], fb)
    with pytest.raises(RuntimeError):
        backends.render_pointmaps("surfelmap", sc["X"], sc["C"], sc["img"], sc["models"], sc["camera"],
                                  sc["proj"], fb)


def test_splat_full_size_properties(backends, tmp_path):
    """512x384 pointmaps, 8 keyframes into 1024x768: deterministic, order-independent, adding a
    keyframe never moves a pixel farther, and the headless window mirror produces the same image."""
    from mast3r_slam.visualization import HeadlessWindow
    from mast3r_slam.sim3 import Sim3

    sc = synth.make_render_scene(K=8, seed=1242, device=DEV)
    w, h = 1024, 768
    for program in ("surfelmap", "trianglemap"):
        a = _gpu_render(backends, program, sc, w, h, radius=0.004)
        b = _gpu_render(backends, program, sc, w, h, radius=0.004)
        assert torch.equal(a, b)
        perm = torch.tensor([5, 2, 7, 0, 3, 6, 1, 4])
        scp = dict(sc, X=sc["X"][perm].contiguous(), C=sc["C"][perm].contiguous(),
                   img=sc["img"][perm].contiguous(), models=sc["models"][perm].contiguous())
        assert torch.equal(a, _gpu_render(backends, program, scp, w, h, radius=0.004))
        sub = dict(sc, X=sc["X"][:4].contiguous(), C=sc["C"][:4].contiguous(), img=sc["img"][:4].contiguous(),
                   models=sc["models"][:4].contiguous())
        s4 = _gpu_render(backends, program, sub, w, h, radius=0.004)
        assert bool((a <= s4).all())  # positive words: smaller = nearer
        _, depth = backends.render_resolve(a)
        assert float((depth < 1).float().mean()) > 0.3
    win = HeadlessWindow(w, h, device=DEV)
    win.pointmap_prog = "trianglemap"
    win.camera, win.proj = sc["camera"], sc["proj"]
    win.render_pointmap(Sim3(sc["T"]), sc["W"], sc["H"], sc["X"], sc["C"], sc["img"])
    rgba, _ = win.read()
    want, _ = backends.render_resolve(_gpu_render(backends, "trianglemap", sc, w, h))
    assert torch.equal(rgba, want)
    win.save_ppm(str(tmp_path / "view.ppm"))
    assert (tmp_path / "view.ppm").stat().st_size == len(b"P6\n1024 768\n255\n") + w * h * 3
