"""`Sim3`: the slice of lietorch.Sim3 the hot path's callers use (reference tracker.py:180,195,212;
geometry.py:46; global_opt.py:115,139; frame.py:24), for installations without lietorch (it is
an un-vendored git dependency of the reference, pyproject.toml:15).  Data layout
[t(3), q(x,y,z,w), s]; composition, inverse, action and left retraction follow the device
helpers the reference itself restates (gn_kernels.cu:177-413).  Poses are 8 floats: these are
O(1) tensor expressions on whatever device the data lives on; bulk work (retraction of all
keyframes, point transforms inside solves) is done by the CUDA kernels."""
import torch

import mast3r_slam_backends as _be


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack((aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz), -1)


def _qrot(q, X):
    v = q[..., :3].expand(X.shape)
    uv = 2.0 * torch.linalg.cross(v, X, dim=-1)
    return X + q[..., 3:4] * uv + torch.linalg.cross(v, uv, dim=-1)


class Sim3:
    embedded_dim = 8
    manifold_dim = 7

    def __init__(self, data):
        self.data = data

    @classmethod
    def Identity(cls, *batch, device="cuda", dtype=torch.float32):
        d = torch.zeros(*batch, 8, device=device, dtype=dtype)
        d[..., 6:8] = 1.0
        return cls(d)

    @property
    def shape(self):
        return self.data.shape[:-1]

    @property
    def device(self):
        return self.data.device

    def __getitem__(self, i):
        return Sim3(self.data[i])

    def inv(self):
        t, q, s = self.data[..., :3], self.data[..., 3:7], self.data[..., 7:]
        qc = torch.cat((-q[..., :3], q[..., 3:]), -1)
        return Sim3(torch.cat((-_qrot(qc, t) / s, qc, 1.0 / s), -1))

    def __mul__(self, other):
        ta, qa, sa = self.data[..., :3], self.data[..., 3:7], self.data[..., 7:]
        tb, qb, sb = other.data[..., :3], other.data[..., 3:7], other.data[..., 7:]
        return Sim3(torch.cat((ta + sa * _qrot(qa, tb), _qmul(qa, qb), sa * sb), -1))

    def act(self, X):
        return self.data[..., 7:] * _qrot(self.data[..., 3:7], X) + self.data[..., :3]

    def retr(self, xi):
        """Exp(xi) * self with xi = [tau, phi, sigma] — m3r_pose_retr (pose_retr_kernel)."""
        out = self.data.reshape(-1, 8).clone().contiguous()
        _be.pose_retr(out, xi.reshape(-1, 7).to(out.dtype).contiguous(), num_fix=0)
        return Sim3(out.view(self.data.shape))

    def matrix(self):
        q = self.data[..., 3:7]
        eye = torch.eye(3, device=q.device, dtype=q.dtype).expand(q.shape[:-1] + (3, 3))
        R = torch.stack([_qrot(q, eye[..., :, k]) for k in range(3)], -1)
        M = torch.zeros(q.shape[:-1] + (4, 4), device=q.device, dtype=q.dtype)
        M[..., :3, :3] = self.data[..., 7:, None] * R
        M[..., :3, 3] = self.data[..., :3]
        M[..., 3, 3] = 1.0
        return M
