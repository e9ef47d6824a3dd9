"""CPU restatement (torch, float32) of the two streaming steps either side of the solvers.
TEST INFRASTRUCTURE ONLY — never imported by the product.

  edge_build       FactorGraph.add_factors, reference global_opt.py:55-67 (per direction)
  add_factors_gate reference global_opt.py:62-87 (match fractions and edge acceptance)
  update_pointmap  Frame.update_pointmap, reference frame.py:41-105 (per-point modes)
Pinned by tests/golden/frontend_ref.npz, produced by the reference's OWN add_factors /
update_pointmap (tests/golden/gen_frontend_golden.py)."""
import numpy as np
import torch


def edge_build(idx_src2dst, valid_match, Q_src, Q_dst, Q_conf):
    batch_inds = torch.arange(idx_src2dst.shape[0])[:, None].repeat(1, idx_src2dst.shape[1])
    # IEEE-correct sqrt (numpy).  torch's CPU sqrt kernel is NOT correctly rounded in this build
    # (e.g. sqrt(32.430938720703125f) -> 5.6948166 instead of 5.6948171); torch's CUDA sqrt — what
    # the reference runs in production — and sqrt.rn.f32 are, and agree with numpy bit for bit.
    Q = torch.from_numpy(np.sqrt((Q_src[batch_inds, idx_src2dst] * Q_dst).numpy()))
    valid = valid_match & (Q > Q_conf)
    return Q, valid.sum(dim=(1, 2))


def add_factors_gate(ii, jj, cnt_j, cnt_i, n, min_match_frac):
    frac = torch.minimum(cnt_j / n, cnt_i / n)
    invalid = frac < min_match_frac
    consecutive = torch.as_tensor(ii) == (torch.as_tensor(jj) - 1)
    return ~((~consecutive) & invalid)


def _to_spherical(P):
    r = torch.linalg.norm(P, dim=-1, keepdim=True)
    x, y, z = torch.tensor_split(P, 3, dim=-1)
    return torch.cat((r, torch.atan2(y, x), torch.acos(z / r)), dim=-1)


def _to_cartesian(s):
    r, phi, theta = torch.tensor_split(s, 3, dim=-1)
    return torch.cat((r * torch.sin(theta) * torch.cos(phi), r * torch.sin(theta) * torch.sin(phi),
                      r * torch.cos(theta)), dim=-1)


def update_pointmap(X_canon, C, X, Cn, mode):
    """-> (X_canon', C') for the per-point modes of frame.py:69-101."""
    if mode == "weighted_pointmap":
        return ((C * X_canon) + (Cn * X)) / (C + Cn), C + Cn
    if mode == "weighted_spherical":
        s = ((C * _to_spherical(X_canon)) + (Cn * _to_spherical(X))) / (C + Cn)
        return _to_cartesian(s), C + Cn
    if mode == "indep_conf":
        m = Cn > C
        Xo, Co = X_canon.clone(), C.clone()
        Xo[m.repeat(1, 3)] = X[m.repeat(1, 3)]
        Co[m] = Cn[m]
        return Xo, Co
    raise ValueError(mode)
