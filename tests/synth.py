"""Seeded synthetic inputs for the hot path (MASt3R weights are not available offline).

Shapes and statistics follow SURVEY.md 8(d): 512x384 pointmaps from a smooth analytic surface
seen through the TUM fr1 pinhole scaled to 512 px, confidences 1+exp(.), 24-d unit descriptors.
Everything is generated with torch on the requested device from a torch.Generator seed, so the
GPU tests, the CPU oracle and bench.py see identical inputs.  Test/bench infrastructure only.
"""
import math

import torch
import torch.nn.functional as F

W0, H0 = 512, 384
K_TUM_512 = (413.84, 413.20, 254.88, 204.24)  # fx fy cx cy  (dataloader.py:78-80 scaled by 512/640)


def intrinsics(H=H0, W=W0, device="cpu"):
    sx, sy = W / W0, H / H0
    fx, fy, cx, cy = K_TUM_512
    K = torch.tensor([[fx * sx, 0.0, cx * sx], [0.0, fy * sy, cy * sy], [0.0, 0.0, 1.0]],
                     dtype=torch.float32, device=device)
    return K


def _gen(seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    return g


def _blur(x, sigma):
    """Separable Gaussian blur of a (B,C,H,W) tensor, reflect-free (replicate) padding."""
    r = max(1, int(3 * sigma))
    t = torch.arange(-r, r + 1, device=x.device, dtype=x.dtype)
    k = torch.exp(-0.5 * (t / sigma) ** 2)
    k = k / k.sum()
    C = x.shape[1]
    x = F.conv2d(F.pad(x, (r, r, 0, 0), mode="replicate"), k.view(1, 1, 1, -1).repeat(C, 1, 1, 1), groups=C)
    x = F.conv2d(F.pad(x, (0, 0, r, r), mode="replicate"), k.view(1, 1, -1, 1).repeat(C, 1, 1, 1), groups=C)
    return x


def smooth_noise(shape_bchw, sigma, gen, device):
    x = torch.randn(shape_bchw, generator=gen, device=device, dtype=torch.float32)
    x = _blur(x, sigma)
    return x / x.std().clamp_min(1e-6)


def pixel_grid(H, W, device):
    v, u = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32),
                          torch.arange(W, device=device, dtype=torch.float32), indexing="ij")
    return u, v


def surface_z(x, y):
    """World surface z = f(x, y): smooth, in front of every camera (depth 1.4 .. 2.8)."""
    return 2.0 + 0.35 * torch.sin(1.7 * x + 0.3) * torch.cos(1.3 * y) + 0.15 * x - 0.1 * y


# ----------------------------------------------------------------------------- Sim3 (torch)

def quat_mul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], -1)


def quat_rot(q, X):
    qv, w = q[..., :3], q[..., 3:4]
    uv = 2.0 * torch.cross(qv.expand_as(X), X, dim=-1)
    return X + w * uv + torch.cross(qv.expand_as(X), uv, dim=-1)


def sim3_act(T, X):
    """T (...,8) [t,q(xyzw),s]; X (...,3) -> s R X + t"""
    return T[..., 7:8] * quat_rot(T[..., 3:7], X) + T[..., 0:3]


def sim3_inv_act(T, P):
    q = T[..., 3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0], device=T.device)
    return quat_rot(q, P - T[..., 0:3]) / T[..., 7:8]


def axis_angle_quat(axis, angle):
    axis = axis / axis.norm(dim=-1, keepdim=True)
    h = 0.5 * angle
    return torch.cat([axis * torch.sin(h)[..., None], torch.cos(h)[..., None]], -1)


# ----------------------------------------------------------------------------- matching

def make_match_inputs(b=1, H=H0, W=W0, Fdim=24, seed=1235, device="cpu", max_disp=12.0):
    """SURVEY 8(d) config 2.  Returns dict(X11, X21, D11, D21 (b,H,W,.) f32, p_star (b,H,W,2))."""
    gen = _gen(seed, device)
    K = intrinsics(H, W, device)
    u, v = pixel_grid(H, W, device)
    z = (2.0 + 0.5 * torch.sin(2 * math.pi * u / W * 1.5) * torch.cos(2 * math.pi * v / H) + 0.3 * u / W)
    z = z[None] + 0.05 * smooth_noise((b, 1, H, W), 8.0 * W / W0, gen, device)[:, 0]
    z = z.clamp(0.8, 4.0)
    X11 = torch.stack([(u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z], -1).contiguous()
    disp = smooth_noise((b, 2, H, W), 32.0 * W / W0, gen, device) * (max_disp * W / W0 / 2.5)
    disp = disp.clamp(-max_disp * W / W0, max_disp * W / W0)
    # clamp bounds are deliberately non-integer: an exactly integer optimum would sit on the
    # .long() truncation boundary of matching.py:68 and make the integer match index ill-posed
    pu = (u[None] + disp[:, 0]).clamp(2.37, W - 3.37)
    pv = (v[None] + disp[:, 1]).clamp(2.37, H - 3.37)
    grid = torch.stack([pu / (W - 1) * 2 - 1, pv / (H - 1) * 2 - 1], -1)
    samp = lambda img: F.grid_sample(img.permute(0, 3, 1, 2), grid, mode="bilinear",
                                     align_corners=True).permute(0, 2, 3, 1)
    X21 = samp(X11) * (1.0 + 2e-3 * torch.randn((b, H, W, 1), generator=gen, device=device))
    D11 = F.normalize(_blur(torch.randn((b, Fdim, H, W), generator=gen, device=device), 3.0 * W / W0)
                      .permute(0, 2, 3, 1), dim=-1).contiguous()
    D21 = F.normalize(samp(D11) + 0.05 * torch.randn((b, H, W, Fdim), generator=gen, device=device), dim=-1)
    return dict(X11=X11.contiguous(), X21=X21.contiguous(), D11=D11, D21=D21.contiguous(),
                p_star=torch.stack([pu, pv], -1), K=K)


# ----------------------------------------------------------------------------- factor graph

def make_edges(N, counts, seed):
    """Undirected edge list: consecutive / skip-k chains plus seeded random long-range edges.
    counts: dict skip -> how many (None = all possible), plus key 'random' -> number."""
    g = torch.Generator().manual_seed(int(seed))
    edges = []
    seen = set()
    for skip, cnt in counts.items():
        if skip == "random":
            continue
        all_e = [(i, i + skip) for i in range(N - skip)]
        if cnt is not None:
            all_e = all_e[:cnt]
        for e in all_e:
            if e not in seen:
                seen.add(e); edges.append(e)
    nrand = counts.get("random", 0)
    tries = 0
    while nrand > 0 and tries < 100000:
        tries += 1
        i = int(torch.randint(0, N, (1,), generator=g)); j = int(torch.randint(0, N, (1,), generator=g))
        if i > j: i, j = j, i
        if j - i < 2 or (i, j) in seen: continue
        seen.add((i, j)); edges.append((i, j)); nrand -= 1
    ii = torch.tensor([e[0] for e in edges], dtype=torch.int64)
    jj = torch.tensor([e[1] for e in edges], dtype=torch.int64)
    return ii, jj


def make_poses(N, seed, device, radius=0.5):
    """GT Sim3 world<-camera poses on an arc facing the surface, scales U(0.9,1.1)."""
    gen = _gen(seed, device)
    ang = torch.linspace(-0.6, 0.6, N, device=device)
    t = torch.stack([radius * torch.sin(ang) * 2.0, 0.15 * torch.cos(3 * ang), radius * (1 - torch.cos(ang))], -1)
    axis = torch.tensor([0.0, 1.0, 0.0], device=device).expand(N, 3) + 0.1 * torch.randn((N, 3), generator=gen, device=device)
    q = axis_angle_quat(axis, -0.25 * ang)  # look slightly inward
    s = 0.9 + 0.2 * torch.rand((N, 1), generator=gen, device=device)
    s[0] = 1.0
    return torch.cat([t, q, s], -1).float().contiguous()


def render_pointmaps(T_wc, H, W, K, iters=10):
    """Pointmap of the analytic surface in each camera's own (scaled) frame: (N,H*W,3)."""
    device = T_wc.device
    N = T_wc.shape[0]
    u, v = pixel_grid(H, W, device)
    d_cam = torch.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)], -1).view(1, -1, 3)
    Tn = T_wc[:, None, :]
    d_w = quat_rot(Tn[..., 3:7], d_cam.expand(N, -1, -1))
    c = Tn[..., 0:3]
    tt = torch.full((N, H * W, 1), 2.0, device=device)
    for _ in range(iters):
        P = c + tt * d_w
        tt = (surface_z(P[..., 0:1], P[..., 1:2]) - c[..., 2:3]) / d_w[..., 2:3]
    Pw = c + tt * d_w
    return sim3_inv_act(Tn, Pw).contiguous()


def perturb_poses(T, seed, sig_t=0.01, sig_r=math.radians(1.0), sig_s=0.01):
    gen = _gen(seed, T.device)
    N = T.shape[0]
    dt = sig_t * torch.randn((N, 3), generator=gen, device=T.device)
    ax = torch.randn((N, 3), generator=gen, device=T.device)
    dq = axis_angle_quat(ax, sig_r * torch.randn((N,), generator=gen, device=T.device))
    ds = torch.exp(sig_s * torch.randn((N, 1), generator=gen, device=T.device))
    out = torch.cat([T[:, 0:3] + dt, quat_mul(dq, T[:, 3:7]), T[:, 7:8] * ds], -1)
    out[0] = T[0]
    return out.float().contiguous()


def edge_streams(T_gt, Xs, ii_d, jj_d, H, W, K, seed, chunk=64, edge_ids=None):
    """Per directed edge (i <- j): idx (M,n) i64 nearest-pixel projection of T_ij Xj into camera i,
    valid (M,n,1) bool (inside image and depth-consistent within 5 %), Q (M,n,1) f32.
    edge_ids: GLOBAL ids of these directed edges.  With them every edge draws its confidence noise
    from its own generator (seed, id) and is normalised on its own, so an edge's streams do not
    depend on which other edges — i.e. which rank's slice — it is generated with: the k-GPU problem
    IS the 1-GPU problem.  Without (default) the noise comes from one
    generator in chunks, as the committed golden fixtures were made."""
    device = Xs.device
    gen = _gen(seed, device)
    M, n = ii_d.shape[0], Xs.shape[1]
    idx = torch.empty((M, n), dtype=torch.int64, device=device)
    valid = torch.empty((M, n, 1), dtype=torch.bool, device=device)
    Q = torch.empty((M, n, 1), dtype=torch.float32, device=device)
    for s in range(0, M, chunk):
        i = ii_d[s:s + chunk].to(device); j = jj_d[s:s + chunk].to(device)
        Pw = sim3_act(T_gt[j][:, None, :], Xs[j])
        Xi = sim3_inv_act(T_gt[i][:, None, :], Pw)
        z = Xi[..., 2].clamp_min(1e-3)
        uu = torch.round(K[0, 0] * Xi[..., 0] / z + K[0, 2])
        vv = torch.round(K[1, 1] * Xi[..., 1] / z + K[1, 2])
        inside = (uu >= 0) & (uu <= W - 1) & (vv >= 0) & (vv <= H - 1) & (Xi[..., 2] > 1e-3)
        lin = (vv.clamp(0, H - 1) * W + uu.clamp(0, W - 1)).long()
        zi = torch.gather(Xs[i][..., 2], 1, lin)
        ok = inside & ((zi - Xi[..., 2]).abs() < 0.05 * Xi[..., 2].abs())
        idx[s:s + chunk] = torch.where(ok, lin, torch.zeros_like(lin))
        valid[s:s + chunk, :, 0] = ok
        B = i.shape[0]
        if edge_ids is None:
            qn = smooth_noise((B, 1, H, W), 4.0 * W / W0, gen, device).view(B, n)
        else:
            x = torch.stack([torch.randn((1, H, W), generator=_gen(int(seed) * 1000003 + int(e), device), device=device)
                             for e in edge_ids[s:s + chunk]])
            x = _blur(x, 4.0 * W / W0).view(B, n)
            qn = x / x.std(dim=1, keepdim=True).clamp_min(1e-6)
        Q[s:s + chunk, :, 0] = 1.0 + torch.exp(0.7 + 0.5 * qn)
    return idx, valid, Q


def make_graph(N=32, counts=None, H=H0, W=W0, seed=1236, device="cpu", kf_ids=None, noise=1e-3):
    """SURVEY 8(d) config 3/4/5.  Returns dict with the tensors gauss_newton_* takes
    (two-way directed edges as prep_two_way_edges builds them, global_opt.py:104-110)."""
    if counts is None:
        counts = {1: None, 2: 30, 3: 29, "random": 10}
    K = intrinsics(H, W, device)
    T_gt = make_poses(N, seed, device)
    Xs = render_pointmaps(T_gt, H, W, K)
    gen = _gen(seed + 7, device)
    Xs = (Xs * (1.0 + noise * torch.randn(Xs.shape[:2] + (1,), generator=gen, device=device))).contiguous()
    n = H * W
    cn = smooth_noise((N, 1, H, W), 4.0 * W / W0, gen, device).view(N, n, 1)
    Cs = (1.0 + torch.exp(0.5 + 0.5 * cn)).contiguous()
    ii_u, jj_u = make_edges(N, counts, seed)
    ii_d = torch.cat([ii_u, jj_u]); jj_d = torch.cat([jj_u, ii_u])  # prep_two_way_edges
    idx, valid, Q = edge_streams(T_gt, Xs, ii_d, jj_d, H, W, K, seed + 13)
    ids = torch.arange(N, dtype=torch.int64) if kf_ids is None else torch.as_tensor(kf_ids, dtype=torch.int64)
    T0 = perturb_poses(T_gt, seed + 29)
    return dict(Twc=T0, T_gt=T_gt, Xs=Xs, Cs=Cs, ii=ids[ii_d].to(device), jj=ids[jj_d].to(device),
                idx=idx, valid=valid, Q=Q, K=K, H=H, W=W, N=N, M=ii_d.shape[0])


def constrain_points_to_ray(Xs, H, W, K):
    """geometry.constrain_points_to_ray (geometry.py:37-42,107-115): X <- z * K^-1 [u,v,1]."""
    u, v = pixel_grid(H, W, Xs.device)
    z = Xs[..., 2:3]
    dirs = torch.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)], -1).view(1, -1, 3)
    return (z * dirs).contiguous()


# ----------------------------------------------------------------------------- tracker

def make_tracker_inputs(H=H0, W=W0, seed=1234, device="cpu", noise=1e-3):
    """Xk = scene, Xf = T_gt^-1 Xk + noise, identity correspondences."""
    gen = _gen(seed, device)
    m = make_match_inputs(1, H, W, 8, seed, device)
    Xk = m["X11"].view(-1, 3).contiguous()
    axis = torch.tensor([1.0, 2.0, 3.0], device=device)
    q = axis_angle_quat(axis, torch.tensor(math.radians(3.0), device=device))
    T_gt = torch.cat([torch.tensor([0.05, -0.02, 0.03], device=device), q,
                      torch.tensor([1.04], device=device)]).float()
    Xf = (sim3_inv_act(T_gt[None], Xk) + noise * torch.randn(Xk.shape, generator=gen, device=device)).contiguous()
    n = Xk.shape[0]
    Qk = (1.0 + torch.exp(0.7 + 0.5 * smooth_noise((1, 1, H, W), 4.0 * W / W0, gen, device).view(n, 1))).contiguous()
    valid = Qk > 1.5
    return dict(Xf=Xf, Xk=Xk, Qk=Qk, valid=valid, T_gt=T_gt)


def make_tracker_calib_inputs(H=H0, W=W0, seed=1234, device="cpu", noise=1e-3):
    """Calibrated twin of make_tracker_inputs (tracker.py:131-154 with use_calib): both pointmaps
    are ray-constrained (X = z K^-1 [u,v,1]); the frame's points are the keyframe's surface seen
    from a pose offset by T_gt, matched by identity correspondences."""
    d = make_tracker_inputs(H, W, seed, device, noise)
    K = intrinsics(H, W, device)
    Xk = constrain_points_to_ray(d["Xk"][None], H, W, K)[0].contiguous()
    Xf = sim3_inv_act(d["T_gt"][None], Xk)
    gen = _gen(seed + 5, device)
    Xf = (Xf * (1.0 + noise * torch.randn((Xf.shape[0], 1), generator=gen, device=device))).contiguous()
    u, v = pixel_grid(H, W, device)
    meas = torch.cat([u.reshape(-1, 1), v.reshape(-1, 1), torch.log(Xk[:, 2:3])], -1).contiguous()
    valid_meas = Xk[:, 2:3] > 1e-6
    meas = torch.where(valid_meas.expand(-1, 3), meas, torch.zeros_like(meas)).contiguous()
    return dict(Xf=Xf, Xk=Xk, Qk=d["Qk"], valid=d["valid"], meas_k=meas, valid_meas_k=valid_meas, K=K,
                T_gt=d["T_gt"], H=H, W=W)


def make_symmetric_match_outputs(E, H=H0, W=W0, seed=77, device="cpu", weak=()):
    """What mast3r_match_symmetric returns for E keyframe pairs (global_opt.py:41-53): smooth
    correspondences with seeded dropouts, descriptor confidences 1+exp(.); edges listed in `weak`
    get mostly-invalid matches so that the match-fraction gate rejects them."""
    gen = _gen(seed, device)
    n = H * W
    out = {}
    for tag in ("i2j", "j2i"):
        u, v = pixel_grid(H, W, device)
        du = 6.0 * smooth_noise((E, 1, H, W), 16.0 * W / W0, gen, device)[:, 0]
        dv = 6.0 * smooth_noise((E, 1, H, W), 16.0 * W / W0, gen, device)[:, 0]
        uu = (u[None] + du).round().clamp(0, W - 1)
        vv = (v[None] + dv).round().clamp(0, H - 1)
        out["idx_" + tag] = (vv * W + uu).long().view(E, n).contiguous()
        keep = torch.rand((E, n, 1), generator=gen, device=device) > 0.2
        for e in weak:
            keep[e] = torch.rand((n, 1), generator=gen, device=device) > 0.97
        out["valid_" + tag] = keep.contiguous()
    for name in ("Qii", "Qjj", "Qji", "Qij"):
        q = smooth_noise((E, 1, H, W), 4.0 * W / W0, gen, device).view(E, n, 1)
        out[name] = (1.0 + torch.exp(0.7 + 0.5 * q)).contiguous()
    return out


# ----------------------------------------------------------------------------- headless viewer

def sim3_matrix(T):
    """(N,8) [t, q(xyzw), s] -> (N,4,4) row-major [sR t; 0 1]  (lietorch Sim3.matrix())."""
    N = T.shape[0]
    M = torch.zeros((N, 4, 4), dtype=torch.float32, device=T.device)
    eye = torch.eye(3, device=T.device).expand(N, 3, 3)
    for c in range(3):
        M[:, :3, c] = T[:, 7:8] * quat_rot(T[:, 3:7], eye[:, :, c])
    M[:, :3, 3] = T[:, 0:3]
    M[:, 3, 3] = 1.0
    return M.contiguous()


def gl_perspective(fov_y_deg, aspect, near, far):
    t = 1.0 / math.tan(math.radians(fov_y_deg) / 2.0)
    return torch.tensor([[t / aspect, 0, 0, 0], [0, t, 0, 0],
                         [0, 0, -(far + near) / (far - near), -2.0 * far * near / (far - near)],
                         [0, 0, -1, 0]], dtype=torch.float32)


def gl_view_of_pose(T):
    """GL view matrix (row-major) of a viewer camera at the CV-convention pose T (8,), scale ignored."""
    eye = torch.eye(3, device=T.device)
    R = torch.stack([quat_rot(T[None, 3:7], eye[None, :, c])[0] for c in range(3)], -1)  # columns
    V = torch.eye(4, device=T.device)
    V[:3, :3] = R.T
    V[:3, 3] = -(R.T @ T[0:3])
    flip = torch.diag(torch.tensor([1.0, -1.0, -1.0, 1.0], device=T.device))
    return (flip @ V).float().contiguous()


def make_render_scene(K=3, H=H0, W=W0, seed=1240, device="cpu"):
    """K keyframes looking at the analytic surface (same generator as the graph configs) with
    colours, confidences (a low-confidence band below the viewer's 1.5 threshold) and a viewer
    camera placed behind keyframe 0, slightly off-axis.  Returns the tensors the viewer uploads
    as textures (visualization.py:150-152: X, C, uimg) plus model/view/projection matrices."""
    Kc = intrinsics(H, W, device)
    T = make_poses(max(K, 2), seed, device)[:K]
    X = render_pointmaps(T, H, W, Kc).view(K, H, W, 3).contiguous()
    gen = _gen(seed + 3, device)
    cn = smooth_noise((K, 1, H, W), 6.0 * W / W0, gen, device)[:, 0]
    C = (1.0 + torch.exp(0.9 + 0.8 * cn)).contiguous()  # ~8 % below 1.5
    img = torch.sigmoid(1.5 * smooth_noise((K, 3, H, W), 3.0 * W / W0, gen, device)).permute(0, 2, 3, 1).contiguous()
    viewer = T[0].clone()
    viewer[0:3] = viewer[0:3] + quat_rot(viewer[None, 3:7], torch.tensor([[0.15, -0.1, -0.6]], device=device))[0]
    return dict(X=X, C=C, img=img, models=sim3_matrix(T), camera=gl_view_of_pose(viewer),
                proj=gl_perspective(55.0, 4.0 / 3.0, 0.05, 50.0), T=T, H=H, W=W, K=K)
