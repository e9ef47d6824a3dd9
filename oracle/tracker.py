"""CPU restatement of the tracker's two-frame Sim3 solve — TEST INFRASTRUCTURE ONLY.

Follows the reference's pure-PyTorch path line by line (paths relative to /root/reference):
  FrameTracker.solve / opt_pose_ray_dist_sim3   mast3r_slam/tracker.py:156-214
  point_to_ray_dist, act_Sim3, skew_sym         mast3r_slam/geometry.py:5-52
  check_convergence, huber                      mast3r_slam/nonlinear_optimizer.py:5-33
The reference manipulates poses through `lietorch.Sim3`, an un-vendored, un-pinned git
dependency (pyproject.toml:15).  Its group arithmetic is restated here from the
reference's own CUDA copy of it (gn_kernels.cu:177-413: data [t, q(xyzw), s]; act = s R X + t;
a*b = (t_a + s_a R_a t_b, q_a q_b, s_a s_b); inv = (-R^T t / s, q*, 1/s); retr(xi) = Exp(xi)*X).

Parity pin: tests/golden/tracker_ref_*.npz hold the outputs of the reference's OWN
tracker.py method bodies, imported from /root/reference with this module's Sim3 injected as
`lietorch` (generator: tests/golden/gen_tracker_golden.py).  lietorch itself is absent, so the
group arithmetic is pinned only against the reference's CUDA restatement: "parity unpinned"
for that part.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import math

import torch

EPS = 1e-6  # gn_kernels.cu:34


def quat_mul(a, b):  # quat_comp, gn_kernels.cu:178-184
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], -1)


def quat_rot(q, X):  # actSO3, gn_kernels.cu:195-205
    qv, w = q[..., :3], q[..., 3:4]
    qv = qv.expand(X.shape)
    uv = 2.0 * torch.linalg.cross(qv, X, dim=-1)
    return X + w * uv + torch.linalg.cross(qv, uv, dim=-1)


def exp_sim3(xi):
    """xi (...,7) [tau, phi, sigma] -> data (...,8).  expSO3 / expSim3, gn_kernels.cu:299-390."""
    tau, phi, sigma = xi[..., 0:3], xi[..., 3:6], xi[..., 6:7]
    scale = torch.exp(sigma)
    theta_sq = (phi * phi).sum(-1, keepdim=True)
    theta = torch.sqrt(theta_sq)
    theta_p4 = theta_sq * theta_sq
    small_t2 = theta_sq < EPS
    th_safe = torch.where(small_t2, torch.ones_like(theta), theta)
    imag = torch.where(small_t2, 0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * theta_p4,
                       torch.sin(0.5 * th_safe) / th_safe)
    real = torch.where(small_t2, 1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * theta_p4,
                       torch.cos(0.5 * th_safe))
    q = torch.cat([imag * phi, real], -1)
    small_s = sigma.abs() < EPS
    small_t = theta.abs() < EPS
    one = torch.ones_like(sigma)
    sg = torch.where(small_s, one, sigma)
    th = torch.where(small_t, one, theta)
    th2 = th * th
    C = torch.where(small_s, one, (scale - 1.0) / sg)
    # sigma ~ 0
    A_s0 = torch.where(small_t, 0.5 * one, (1.0 - torch.cos(th)) / th2)
    B_s0 = torch.where(small_t, one / 6.0, (th - torch.sin(th)) / (th2 * th))
    # sigma != 0
    sg2 = sg * sg
    A_t0 = ((sg - 1.0) * scale + 1.0) / sg2
    B_t0 = (scale * 0.5 * sg2 + scale - 1.0 - sg * scale) / (sg2 * sg)
    a = scale * torch.sin(th)
    b = scale * torch.cos(th)
    c = th2 + sg2
    A_g = (a * sg + (1.0 - b) * th) / (th * c)
    B_g = (C - ((b - 1.0) * sg + a * th) / c) / th2
    A = torch.where(small_s, A_s0, torch.where(small_t, A_t0, A_g))
    B = torch.where(small_s, B_s0, torch.where(small_t, B_t0, B_g))
    c1 = torch.linalg.cross(phi, tau, dim=-1)
    c2 = torch.linalg.cross(phi, c1, dim=-1)
    t = C * tau + A * c1 + B * c2
    return torch.cat([t, q, scale], -1)


class Sim3:
    """Minimal stand-in for lietorch.Sim3 (the members the reference's path uses:
    tracker.py:180,195,212; geometry.py:46; global_opt.py:115,139)."""

    embedded_dim = 8
    manifold_dim = 7

    def __init__(self, data):
        self.data = data

    @staticmethod
    def Identity(*batch, device="cpu", dtype=torch.float32):
        d = torch.zeros(*batch, 8, device=device, dtype=dtype)
        d[..., 6] = 1.0
        d[..., 7] = 1.0
        return Sim3(d)

    @property
    def shape(self):
        return self.data.shape[:-1]

    @property
    def device(self):
        return self.data.device

    def __getitem__(self, i):
        return Sim3(self.data[i])

    def inv(self):  # quat_inv + gn_kernels.cu:252-272 with T_j = identity
        t, q, s = self.data[..., 0:3], self.data[..., 3:7], self.data[..., 7:8]
        qi = q * torch.tensor([-1.0, -1.0, -1.0, 1.0], device=q.device, dtype=q.dtype)
        return Sim3(torch.cat([-quat_rot(qi, t) / s, qi, 1.0 / s], -1))

    def __mul__(self, o):
        ta, qa, sa = self.data[..., 0:3], self.data[..., 3:7], self.data[..., 7:8]
        tb, qb, sb = o.data[..., 0:3], o.data[..., 3:7], o.data[..., 7:8]
        return Sim3(torch.cat([ta + sa * quat_rot(qa, tb), quat_mul(qa, qb), sa * sb], -1))

    def act(self, X):  # actSim3, gn_kernels.cu:207-219
        t, q, s = self.data[..., 0:3], self.data[..., 3:7], self.data[..., 7:8]
        return s * quat_rot(q, X) + t

    def retr(self, xi):  # retrSim3, gn_kernels.cu:392-413
        return Sim3(exp_sim3(xi.to(self.data.dtype).view(self.data.shape[:-1] + (7,)))) * self

    def matrix(self):
        t, q, s = self.data[..., 0:3], self.data[..., 3:7], self.data[..., 7:8]
        eye = torch.eye(3, device=q.device, dtype=q.dtype).expand(q.shape[:-1] + (3, 3))
        R = torch.stack([quat_rot(q, eye[..., :, k]) for k in range(3)], -1)
        M = torch.zeros(q.shape[:-1] + (4, 4), device=q.device, dtype=q.dtype)
        M[..., :3, :3] = s[..., None] * R
        M[..., :3, 3] = t
        M[..., 3, 3] = 1.0
        return M


def skew_sym(x):  # geometry.py:5-9
    b = x.shape[:-1]
    a, bb, c = x.unbind(-1)
    o = torch.zeros_like(a)
    return torch.stack([o, -c, bb, c, o, -a, -bb, a, o], -1).view(*b, 3, 3)


def point_to_ray_dist(X, jacobian=False):  # geometry.py:17-34
    d = torch.linalg.norm(X, dim=-1, keepdim=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = torch.cat((r, d), -1)
    if not jacobian:
        return rd
    eye = torch.eye(3, device=X.device, dtype=X.dtype).expand(X.shape[:-1] + (3, 3))
    dr_dX = d_inv.unsqueeze(-1) * (eye - (d_inv ** 2).unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
    return rd, torch.cat((dr_dX, r.unsqueeze(-2)), -2)


def act_sim3_jac(T, pC):  # geometry.py:45-52
    pW = T.act(pC)
    eye = torch.eye(3, device=pW.device, dtype=pW.dtype).expand(pW.shape[:-1] + (3, 3))
    return pW, torch.cat([eye, -skew_sym(pW), pW.unsqueeze(-1)], -1)


def huber(r, k=1.345):  # nonlinear_optimizer.py:28-33
    r_abs = r.abs()
    return torch.where(r_abs < k, torch.ones_like(r), k / r_abs)


def solve(sqrt_info, r, J, huber_k):  # tracker.py:156-171
    robust = sqrt_info * torch.sqrt(huber(sqrt_info * r, huber_k))
    A = (robust[..., None] * J).reshape(-1, J.shape[-1])
    b = (robust * r).reshape(-1, 1)
    H = A.T @ A
    g = -A.T @ b
    cost = 0.5 * float(b.T @ b)
    L = torch.linalg.cholesky(H, upper=False)  # raises on failure (caught at tracker.py:91)
    return torch.cholesky_solve(g, L, upper=False).view(1, -1), cost, H, g


def opt_pose_ray_dist_sim3(Xf, Xk, T_CkCf, Qk, valid, sigma_ray=0.003, sigma_dist=10.0,
                           huber_k=1.345, max_iters=50, rel_error=1e-3, delta_norm=1e-3,
                           history=None):
    """tracker.py:173-214 on the RELATIVE pose (the `T_WCk.inv() * T_WCf` / `T_WCk * T_CkCf`
    bracketing at :180,:212 is caller arithmetic).  T_CkCf: Sim3 with data (1,8).
    -> (T_CkCf, iterations, converged).  `history` (list) receives (cost, tau, data) per step."""
    sqrt_info_ray = 1.0 / sigma_ray * valid * torch.sqrt(Qk)
    sqrt_info_dist = 1.0 / sigma_dist * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_ray.repeat(1, 3), sqrt_info_dist), 1)
    rd_k = point_to_ray_dist(Xk)
    old_cost = float("inf")
    its, conv = 0, False
    for step in range(max_iters):
        Xf_Ck, dX_dT = act_sim3_jac(T_CkCf, Xf)
        rd_f, drd_dX = point_to_ray_dist(Xf_Ck, jacobian=True)
        r = rd_k - rd_f
        J = -drd_dX @ dX_dT
        tau, new_cost, _, _ = solve(sqrt_info, r, J, huber_k)
        T_CkCf = T_CkCf.retr(tau)
        its = step + 1
        if history is not None:
            history.append((new_cost, tau.clone(), T_CkCf.data.clone()))
        # check_convergence, nonlinear_optimizer.py:5-26
        rel_dec = math.fabs((old_cost - new_cost) / old_cost)
        if rel_dec < rel_error or float(torch.linalg.norm(tau)) < delta_norm:
            conv = True
            break
        old_cost = new_cost
    return T_CkCf, its, conv


def project_calib(P, K, img_size, border=0, z_eps=0.0):  # geometry.py:63-104
    """-> pz (..,3) = [u, v, log z], dpz_dP (..,3,3), valid (..,1)."""
    b = P.shape[:-1]
    p = (K.repeat(*b, 1, 1) @ P[..., None]).squeeze(-1)
    p = p / p[..., 2:3]
    p = p[..., :2]
    u, v = p.split([1, 1], dim=-1)
    x, y, z = P.split([1, 1, 1], dim=-1)
    valid_u = (u > border) & (u < img_size[1] - 1 - border)
    valid_v = (v > border) & (v < img_size[0] - 1 - border)
    valid_z = z > z_eps
    valid = valid_u & valid_v & valid_z
    logz = torch.log(z)
    logz[torch.logical_not(valid_z)] = 0.0
    pz = torch.cat((p, logz), dim=-1)
    fx, fy = K[..., 0, 0], K[..., 1, 1]
    z_inv = 1.0 / z[..., 0]
    J = torch.zeros(*b + (3, 3), device=P.device, dtype=P.dtype)
    J[..., 0, 0] = fx
    J[..., 1, 1] = fy
    J[..., 0, 2] = -fx * x[..., 0] * z_inv
    J[..., 1, 2] = -fy * y[..., 0] * z_inv
    J *= z_inv[..., None, None]
    J[..., 2, 2] = z_inv
    return pz, J, valid


def calib_measurements(Xk, img_size, depth_eps):  # tracker.py:143-152 (get_points_poses)
    """Xk (n,3) already ray-constrained -> meas_k (n,3) = [u, v, log z_k], valid_meas_k (n,1)."""
    h, w = img_size
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")
    uv = torch.stack((u, v), dim=-1).to(Xk.dtype).view(-1, 2)
    meas = torch.cat((uv, torch.log(Xk[..., 2:3])), dim=-1)
    valid = Xk[..., 2:3] > depth_eps
    meas[~valid.repeat(1, 3)] = 0.0
    return meas, valid


def opt_pose_calib_sim3(Xf, T_CkCf, Qk, valid, meas_k, valid_meas_k, K, img_size, sigma_pixel=1.0,
                        sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6, huber_k=1.345,
                        max_iters=50, rel_error=1e-3, delta_norm=1e-3, history=None):
    """tracker.py:216-266 on the RELATIVE pose.  -> (T_CkCf, iterations, converged)."""
    sqrt_info_pixel = 1.0 / sigma_pixel * valid * torch.sqrt(Qk)
    sqrt_info_depth = 1.0 / sigma_depth * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((sqrt_info_pixel.repeat(1, 2), sqrt_info_depth), 1)
    old_cost = float("inf")
    its, conv = 0, False
    for step in range(max_iters):
        Xf_Ck, dX_dT = act_sim3_jac(T_CkCf, Xf)
        pz, dpz_dX, valid_proj = project_calib(Xf_Ck, K, img_size, pixel_border, depth_eps)
        sqrt_info2 = (valid_proj & valid_meas_k) * sqrt_info
        r = meas_k - pz
        J = -dpz_dX @ dX_dT
        tau, new_cost, _, _ = solve(sqrt_info2, r, J, huber_k)
        T_CkCf = T_CkCf.retr(tau)
        its = step + 1
        if history is not None:
            history.append((new_cost, tau.clone(), T_CkCf.data.clone()))
        rel_dec = math.fabs((old_cost - new_cost) / old_cost)
        if rel_dec < rel_error or float(torch.linalg.norm(tau)) < delta_norm:
            conv = True
            break
        old_cost = new_cost
    return T_CkCf, its, conv
