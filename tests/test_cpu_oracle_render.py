"""Oracle of the headless viewer splat (oracle/m3r_render_oracle.c): known-answer tests on the
shader restatement.  Parity is UNPINNED (no OpenGL anywhere to run the reference's GLSL): these
tests pin the oracle to closed forms of the shader source instead (surfelmap.glsl / trianglemap.glsl).
"""
import math

import numpy as np

from oracle import render as R

CLEAR = 0xFF26231E


def _plane(H, W, f, z0, slope=0.0):
    u, v = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    z = (z0 + slope * (u - W / 2) / f).astype(np.float32)
    X = np.stack([(u - W / 2) / f * z, (v - H / 2) / f * z, z], -1)[None].astype(np.float32)
    return X


def _gl(H, W, f, scale):
    cam = np.diag([1, -1, -1, 1]).astype(np.float32)  # CV camera at the origin -> GL view
    n, fa = 0.1, 100.0
    t = 2 * f / H
    proj = np.array([[2 * f / W, 0, 0, 0], [0, t, 0, 0], [0, 0, -(fa + n) / (fa - n), -2 * fa * n / (fa - n)],
                     [0, 0, -1, 0]], np.float32)
    return cam, proj, scale * W, scale * H


def test_surfel_disk_geometry_and_depth():
    H, W, f, z0 = 8, 8, 10.0, 2.0
    X = _plane(H, W, f, z0)
    C = np.zeros((1, H, W), np.float32)
    C[0, 4, 4] = 2.0  # a single confident surfel at the principal point
    img = np.full((1, H, W, 3), 0.5, np.float32)
    cam, proj, w, h = _gl(H, W, f, 16)
    r = 0.05
    fb = R.render("surfelmap", X, C, img, np.eye(4, dtype=np.float32)[None], cam, proj, w, h, radius=r)
    rgba, depth = R.resolve(fb)
    m = depth < 1.0
    # disk of radius r at depth z0 seen through a focal length of 16 f pixels
    rad_px = r / z0 * f * 16
    assert abs(m.sum() - math.pi * rad_px ** 2) < 0.08 * math.pi * rad_px ** 2
    ys, xs = np.nonzero(m)
    assert abs(xs.mean() + 0.5 - w / 2) < 0.51 and abs(ys.mean() + 0.5 - h / 2) < 0.51
    # window depth of view-space z = -z0 under the projection
    n, fa = 0.1, 100.0
    zndc = ((fa + n) / (fa - n) * z0 - 2 * fa * n / (fa - n)) / z0
    assert np.allclose(depth[m], 0.5 * zndc + 0.5, atol=2e-6)
    assert (rgba[~m] == [0x1E, 0x23, 0x26, 0xFF]).all()
    # Phong terms at the disk centre: N = (0,0,1) in view space, position (0,0,-z0)
    L = np.array([0.1, 0.1, 0.0]) - np.array([0, 0, -z0])
    L /= np.linalg.norm(L)
    lam = L[2]
    Rv = 2 * (L * [0, 0, 1]) * [0, 0, 1] - L  # component-wise product, as in the shader
    spec = max(Rv @ [0, 0, 1], 0) ** 32
    want = 0.5 * (0.3 + lam * 0.5) + 0.4 * spec
    centre = rgba[h // 2, w // 2, :3].astype(np.float64) / 255
    assert np.allclose(centre, want, atol=2 / 255)


def test_surfel_gates_border_conf_cull():
    H, W, f = 8, 8, 10.0
    X = _plane(H, W, f, 2.0)
    img = np.full((1, H, W, 3), 0.5, np.float32)
    cam, proj, w, h = _gl(H, W, f, 8)
    I = np.eye(4, dtype=np.float32)[None]
    C = np.zeros((1, H, W), np.float32)
    C[0, 0, :] = C[0, -1, :] = C[0, :, 0] = C[0, :, -1] = 9.0  # border pixels only: never drawn (:32-34)
    assert (R.resolve(R.render("surfelmap", X, C, img, I, cam, proj, w, h, radius=0.05))[1] == 1.0).all()
    C[:] = 1.49  # below conf_threshold everywhere (:36-39); >= threshold draws
    assert (R.resolve(R.render("surfelmap", X, C, img, I, cam, proj, w, h, radius=0.05))[1] == 1.0).all()
    C[:] = 1.5
    front = R.resolve(R.render("surfelmap", X, C, img, I, cam, proj, w, h, radius=0.05))[1] < 1.0
    assert front.sum() > 0
    # seen from behind (viewer rotated 180 deg about y and placed beyond the plane) the quads are back faces
    cam_b = np.array([[-1, 0, 0, 0], [0, -1, 0, 0], [0, 0, 1, -4.0], [0, 0, 0, 1]], np.float32)
    assert (R.resolve(R.render("surfelmap", X, C, img, I, cam_b, proj, w, h, radius=0.05))[1] == 1.0).all()
    back = R.resolve(R.render("surfelmap", X, C, img, I, cam_b, proj, w, h, radius=0.05, cull_face=0))[1] < 1.0
    assert back.sum() == front.sum()


def test_depth_test_keeps_nearest_in_any_order():
    H, W, f = 8, 8, 10.0
    near, far = _plane(H, W, f, 1.5), _plane(H, W, f, 3.0)
    C = np.full((2, H, W), 2.0, np.float32)
    img = np.stack([np.full((H, W, 3), 0.9, np.float32), np.full((H, W, 3), 0.1, np.float32)])
    cam, proj, w, h = _gl(H, W, f, 8)
    I2 = np.stack([np.eye(4, dtype=np.float32)] * 2)
    a = R.render("surfelmap", np.concatenate([near, far]), C, img, I2, cam, proj, w, h, radius=0.2)
    b = R.render("surfelmap", np.concatenate([far, near]), C, img[::-1].copy(), I2, cam, proj, w, h, radius=0.2)
    assert np.array_equal(a, b)
    only_near = R.render("surfelmap", near, C[:1], img[:1], I2[:1], cam, proj, w, h, radius=0.2)
    covered = R.resolve(only_near)[1] < 1.0
    assert np.array_equal(R.resolve(a)[0][covered], R.resolve(only_near)[0][covered])


def test_trianglemap_watertight_slant_and_colour_interpolation():
    H, W, f = 32, 40, 30.0
    X = _plane(H, W, f, 2.0)
    C = np.full((1, H, W), 2.0, np.float32)
    u = np.arange(W, dtype=np.float32)[None, :, None] / W
    img = np.broadcast_to(u, (H, W, 1)).repeat(3, -1)[None].astype(np.float32).copy()
    cam, proj, w, h = _gl(H, W, f, 4)
    I = np.eye(4, dtype=np.float32)[None]
    fb = R.render("trianglemap", X, C, img, I, cam, proj, w, h, show_normal=0)
    rgba, depth = R.resolve(fb)
    m = depth < 1.0
    # quads x,y in [10, W-10) x [10, H-10) cover pointmap pixels 10 .. W-10 / H-10: a solid rectangle
    # whose window-space edges are pixel centres 10 and W-10 (H-10), 4 target pixels per pointmap pixel
    assert m.sum() == (4 * (W - 20)) * (4 * (H - 20))
    ys, xs = np.nonzero(m)
    assert xs.min() == 40 and xs.max() == 4 * (W - 10) - 1 and ys.min() == 40 and ys.max() == 4 * (H - 10) - 1
    # show_normal output of a fronto-parallel plane: N_view = (0,0,1) -> (-N.x, N.y, N.z)*0.5+0.5
    nrm = R.resolve(R.render("trianglemap", X, C, img, I, cam, proj, w, h, show_normal=1))[0]
    assert (nrm[m][:, :3] == [128, 128, 255]).all()
    # ambient-only shading: the colour ramps linearly with the pointmap column (perspective-correct
    # interpolation on a fronto-parallel plane is linear)
    flat = R.resolve(R.render("trianglemap", X, C, img, I, cam, proj, w, h, phong=(1.0, 0.0, 0.0)))[0]
    row = flat[h // 2, :, 0].astype(np.float64)
    cols = np.arange(44, 4 * (W - 10) - 4)
    lin = np.polyfit(cols, row[cols], 1)
    assert np.abs(np.polyval(lin, cols) - row[cols]).max() <= 1.0
    assert abs(lin[0] - 255.0 / (4 * W)) < 0.02
    # the slant test (:65-70) drops quads whose normal is nearly perpendicular to the viewing ray:
    # plane z = 2 + 12 x under a narrow field of view, |n . ray| = |1 - 12 a| / (|ray| sqrt(145))
    f2 = 300.0
    u2, v2 = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    a = (u2 - W / 2) / f2
    z = 2.0 / (1.0 - 12.0 * a)
    Xs = np.stack([a * z, (v2 - H / 2) / f2 * z, z], -1)[None].astype(np.float32)
    cam2, proj2, w2, h2 = _gl(H, W, f2, 4)
    keep = R.resolve(R.render("trianglemap", Xs, C, img, I, cam2, proj2, w2, h2))[1] < 1.0
    every = R.resolve(R.render("trianglemap", Xs, C, img, I, cam2, proj2, w2, h2, slant_threshold=0.0))[1] < 1.0
    none = R.resolve(R.render("trianglemap", Xs, C, img, I, cam2, proj2, w2, h2, slant_threshold=0.99))[1] < 1.0
    assert every.sum() == (4 * (W - 20)) * (4 * (H - 20)) and none.sum() == 0
    assert (keep <= every).all() and 0 < keep.sum() < every.sum()
    # kept quads are those with 1 - 12 a > 0.1 * sqrt(145) (|ray| ~ 1): columns left of a = -0.017
    col_keep = np.nonzero(keep.any(0))[0]
    a_edge = (col_keep.max() / 4.0 - W / 2) / f2
    assert abs((1 - 12 * a_edge) / math.sqrt(145) - 0.1) < 0.01


def test_depth_bias_and_model_matrix():
    H, W, f = 32, 40, 30.0
    X = _plane(H, W, f, 2.0)
    C = np.full((1, H, W), 2.0, np.float32)
    img = np.full((1, H, W, 3), 0.5, np.float32)
    cam, proj, w, h = _gl(H, W, f, 2)
    I = np.eye(4, dtype=np.float32)[None]
    d0 = R.resolve(R.render("trianglemap", X, C, img, I, cam, proj, w, h))[1]
    d1 = R.resolve(R.render("trianglemap", X, C, img, I, cam, proj, w, h, depth_bias=0.001))[1]
    m = d0 < 1.0
    # clip z -= bias  =>  window depth -= 0.5 * bias / w_clip, w_clip = 2
    assert np.allclose(d0[m] - d1[m], 0.5 * 0.001 / 2.0, atol=1e-6)
    # a Sim3 model matrix (scale 2) pushes the surface to depth 4: same footprint, larger depth
    M = np.diag([2, 2, 2, 1]).astype(np.float32)[None]
    d2 = R.resolve(R.render("trianglemap", X, C, img, M, cam, proj, w, h))[1]
    assert np.array_equal(d2 < 1.0, m) and (d2[m] > d0[m]).all()
