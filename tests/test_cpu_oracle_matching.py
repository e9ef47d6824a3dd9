"""CPU tests of the oracle's matching half (oracle/m3r_oracle.c) — no GPU needed.

Pins: (1) golden vectors produced by the UNMODIFIED reference extension on a B200
(tests/golden/ref_matching.npz, generator tests/golden/gen_ref_golden.py); (2) independent
numpy restatements; (3) known-answer cases."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_half_conversions_exhaustive(orc):
    bits = np.arange(65536, dtype=np.uint16)
    ours = orc.h2f(bits)
    ref = bits.view(np.float16).astype(np.float32)
    nan = np.isnan(ref)
    assert np.array_equal(np.isnan(ours), nan)
    assert np.array_equal(ours[~nan].view(np.uint32), ref[~nan].view(np.uint32))
    # every half value round-trips; random floats and every rounding-boundary neighbourhood
    # round like numpy (IEEE round-to-nearest-even == c10::Half / __float2half_rn)
    assert np.array_equal(orc.f2h(ref[~nan]), bits[~nan])
    rng = np.random.default_rng(0)
    x = np.concatenate([
        rng.standard_normal(200000).astype(np.float32) * np.float32(10.0) ** rng.integers(-9, 6, 200000),
        (ref[~nan][:, None].astype(np.float64) * (1 + np.array([-2.0 ** -12, 2.0 ** -12, 2.0 ** -11 * 0.5,
                                                                 -2.0 ** -11 * 0.5]))).astype(np.float32).ravel(),
        np.array([65504.0, 65519.9, 65520.0, 1e6, 5.96e-8, 2.98e-8, 2.9802322e-8, 2.99e-8, 0.0, -0.0],
                 np.float32)]).astype(np.float32)
    with np.errstate(over="ignore"):
        exp = x.astype(np.float16).view(np.uint16)
    assert np.array_equal(orc.f2h(x), exp)


def _refine_numpy(D11, D21, p1, radius, dil):
    """Independent restatement with numpy float16 scalars (each op rounds to fp16)."""
    b, H, W, F = D11.shape
    out = p1.copy()
    for bi in range(b):
        for k in range(p1.shape[1]):
            u0, v0 = int(p1[bi, k, 0]), int(p1[bi, k, 1])
            best, un, vn = np.float16(0), u0, v0
            for d in range(dil, 0, -1):
                for i in range(0, 2 * radius * d + 1, d):
                    for j in range(0, 2 * radius * d + 1, d):
                        u, v = u0 - radius * d + i, v0 - radius * d + j
                        if 0 <= u < W and 0 <= v < H:
                            s = np.float16(0)
                            for t in range(F):
                                s = np.float16(s + np.float16(D21[bi, k, t] * D11[bi, v, u, t]))
                            if s > best:
                                best, un, vn = s, u, v
                u0, v0 = un, vn
            out[bi, k] = (un, vn)
    return out


def test_refine_matches_vs_numpy_float16(orc):
    rng = np.random.default_rng(5)
    b, H, W, F = 1, 14, 18, 8
    D11 = rng.standard_normal((b, H, W, F)).astype(np.float16)
    D21 = rng.standard_normal((b, 40, F)).astype(np.float16)
    p1 = np.stack([rng.integers(0, W, (b, 40)), rng.integers(0, H, (b, 40))], -1).astype(np.int64)
    with np.errstate(over="ignore"):
        exp = _refine_numpy(D11, D21, p1, 2, 3)
    assert np.array_equal(orc.refine_matches(D11, D21, p1, 2, 3), exp)


def test_refine_matches_kats(orc):
    H, W, F = 32, 32, 24
    # H2: all scores <= 0 => unchanged
    D11 = np.abs(np.random.default_rng(1).standard_normal((1, H, W, F))).astype(np.float16)
    D21 = -np.abs(np.random.default_rng(2).standard_normal((1, 5, F))).astype(np.float16)
    p1 = np.array([[[3, 4], [0, 0], [31, 31], [16, 2], [7, 30]]], np.int64)
    assert np.array_equal(orc.refine_matches(D11, D21, p1, 3, 5), p1)
    # planted peak at +5 px in both axes (reachable at dilation 5)
    D11 = np.zeros((1, H, W, F), np.float16); D11[..., 0] = 0.01; D11[0, 21, 15, 1] = 1.0
    D21 = np.zeros((1, 1, F), np.float16); D21[0, 0, 1] = 1.0; D21[0, 0, 0] = 0.5
    assert orc.refine_matches(D11, D21, np.array([[[10, 16]]], np.int64), 3, 5).tolist() == [[[15, 21]]]
    # ties: strict '>' keeps the first in-image candidate in (u outer, v inner) order
    D11 = np.full((1, H, W, F), 0.25, np.float16); D21 = np.full((1, 1, F), 0.25, np.float16)
    out = orc.refine_matches(D11, D21, np.array([[[16, 16]]], np.int64), 1, 1)
    assert out.tolist() == [[[15, 15]]]
    # fp32 instantiation starts from FLT_MIN: zero scores never win
    z = orc.refine_matches(np.zeros((1, H, W, F), np.float32), np.ones((1, 1, F), np.float32),
                           np.array([[[9, 9]]], np.int64), 3, 5)
    assert z.tolist() == [[[9, 9]]]
    # empty n
    assert orc.refine_matches(D11, D21[:, :0], np.zeros((1, 0, 2), np.int64), 3, 5).shape == (1, 0, 2)


def test_refine_matches_golden_from_reference(orc):
    z = np.load(os.path.join(GOLD, "ref_matching.npz"))
    assert np.array_equal(orc.refine_matches(z["D11h"], z["D21h"], z["p1"], 3, 5), z["refine_f16"])
    assert np.array_equal(orc.refine_matches(z["D11h"], z["D21h"], z["p1"], 2, 3), z["refine_f16_r2d3"])
    D11f, D21f = orc.h2f(z["D11h"]), orc.h2f(z["D21h"])  # not the fp32 inputs of the golden: shape check only
    assert orc.refine_matches(D11f, D21f, z["p1"], 3, 5).shape == z["refine_f32"].shape
    assert (z["refine_f16"] != z["p1"]).any()


def test_iter_proj_golden_from_reference(orc):
    z = np.load(os.path.join(GOLD, "ref_matching.npz"))
    p, c = orc.iter_proj(z["img"], z["pts"], z["p0"], 10, 1e-8, 1e-6)
    d = np.abs(p - z["p_new"]).max(-1)
    n = d.size
    # the oracle has no FMA contraction: to-rounding agreement; LM accept/reject ties may flip
    assert (d > 1e-4).sum() <= max(2, n // 1000), (int((d > 1e-4).sum()), float(d.max()))
    assert (c != z["conv"]).sum() <= max(2, n // 1000)
    assert np.median(d) < 1e-5


def _iter_proj_f64(img, pts, p0, max_iter, lam0, thr):
    """Independent float64 LM restatement (same algorithm, no fp32 quirks)."""
    b, H, W, _ = img.shape
    img = img.astype(np.float64)
    out = np.zeros(p0.shape, np.float64)
    conv = np.zeros(p0.shape[:2], bool)

    def samp(bi, u, v):
        u0, v0 = int(np.floor(u)), int(np.floor(v))
        du, dv = u - u0, v - v0
        return ((1 - du) * (1 - dv) * img[bi, v0, u0] + du * (1 - dv) * img[bi, v0, u0 + 1]
                + (1 - du) * dv * img[bi, v0 + 1, u0] + du * dv * img[bi, v0 + 1, u0 + 1])

    for bi in range(b):
        for k in range(p0.shape[1]):
            u = min(max(float(p0[bi, k, 0]), 1.0), W - 2.0)
            v = min(max(float(p0[bi, k, 1]), 1.0), H - 2.0)
            lam, cv = float(lam0), False
            for _ in range(max_iter):
                s = samp(bi, u, v)
                r = s[:3] / np.linalg.norm(s[:3])
                e = r - pts[bi, k]
                cost = e @ e
                gx, gy = s[3:6], s[6:9]
                A = np.array([[gx @ gx + lam, gx @ gy], [gx @ gy, gy @ gy + lam]])
                dlt = np.linalg.solve(A, -np.array([e @ gx, e @ gy]))
                un = min(max(u + dlt[0], 1.0), W - 2.0)
                vn = min(max(v + dlt[1], 1.0), H - 2.0)
                s2 = samp(bi, un, vn)[:3]
                e2 = s2 / np.linalg.norm(s2) - pts[bi, k]
                nc = e2 @ e2
                if nc < cost:
                    u, v, lam, cv = un, vn, lam * 0.1, nc < thr
                else:
                    lam, cv = lam * 10.0, cost < thr
            out[bi, k] = (u, v)
            conv[bi, k] = cv
    return out, conv


def test_iter_proj_vs_float64_numpy(orc):
    from tests import synth
    from oracle import matching as om

    m = synth.make_match_inputs(1, 24, 32, 8, seed=3, device="cpu", max_disp=3.0)
    img, pts, p0 = om.prep_for_iter_proj(m["X11"], m["X21"])
    img, pts, p0 = img.numpy(), pts.numpy(), p0.numpy()
    p, c = orc.iter_proj(img, pts, p0, 10, 1e-8, 1e-6)
    p64, c64 = _iter_proj_f64(img, pts, p0, 10, 1e-8, 1e-6)
    d = np.abs(p - p64).max(-1)
    # fp32 noise floor of the cost (~1e-14 on unit rays) decides late accept/reject steps, so
    # compare where both converged and allow a small tail
    both = c & c64
    assert both.mean() > 0.9
    assert np.median(d[both]) < 2e-3 and (d[both] > 0.05).mean() < 0.01


def test_iter_proj_kats(orc):
    from tests import synth
    from oracle import matching as om

    H, W = 24, 32
    m = synth.make_match_inputs(1, H, W, 8, seed=7, device="cpu")
    img, _, p0 = om.prep_for_iter_proj(m["X11"], m["X11"])
    img, p0 = img.numpy(), p0.numpy()
    # identity warp: stays put (to a few ulps) and converges
    p, c = orc.iter_proj(img, img[..., :3].reshape(1, H * W, 3), p0, 10, 1e-8, 1e-6)
    inner = (p0[..., 0] >= 1) & (p0[..., 0] <= W - 2) & (p0[..., 1] >= 1) & (p0[..., 1] <= H - 2)
    assert np.abs(p - p0)[inner].max() < 1e-4 and c[inner].all()
    # max_iter = 0: clamped init, nothing converged
    p, c = orc.iter_proj(img, img[..., :3].reshape(1, H * W, 3), p0 - 5.0, 0, 1e-8, 1e-6)
    assert np.array_equal(p[..., 0], np.clip(p0[..., 0] - 5, 1, W - 2)) and not c.any()


def test_img_gradient_restatement_vs_reference_golden():
    """oracle.matching.img_gradient against the reference's own image.img_gradient
    (tests/golden/gen_image_golden.py), including a 2x2 image (reflect padding at its minimum)."""
    import torch

    from oracle import matching as om

    z = np.load(os.path.join(GOLD, "image_ref.npz"))
    for name in ("a", "b", "c"):
        gx, gy = om.img_gradient(torch.from_numpy(z[f"{name}_img"]))
        assert np.allclose(gx.numpy(), z[f"{name}_gx"], rtol=0, atol=1e-6)
        assert np.allclose(gy.numpy(), z[f"{name}_gy"], rtol=0, atol=1e-6)
