"""FactorGraph.solve_GN_rays / solve_GN_calib with the reference's behaviour (reference
mast3r_slam/global_opt.py:101-213) on top of mast3r_slam_backends.gauss_newton_{rays,calib}.
Edge construction (`add_factors`, :30-99): the MASt3R network + symmetric matching is caller-side
(`matcher`, the reference's mast3r_match_symmetric); everything after it — Q = sqrt(Q_src[idx] * Q_dst),
the Q_conf gate, match fractions, edge acceptance and appending — is here, the per-point part as
one CUDA launch per direction (m3r_edge_build).  `add_edges` appends already-built tensors.

`frames` is any keyframe store with the reference's protocol: `frames[i]` has `.X_canon (n,3)`,
`.T_WC` (Sim3 with data (1,8)), `.get_average_conf() (n,1)`, `.img (3,h,w)`; and
`frames.update_T_WCs(T_WCs, idx)` (frame.py:309-317)."""
import torch

import mast3r_slam_backends as _be
from mast3r_slam.config import config
from mast3r_slam.geometry import constrain_points_to_ray
from mast3r_slam.sim3 import Sim3


class FactorGraph:
    def __init__(self, model=None, frames=None, K=None, device="cuda"):
        self.model = model
        self.frames = frames
        self.device = device
        self.cfg = config["local_opt"]
        self.K = K
        e = lambda dt: torch.as_tensor([], dtype=dt, device=device)
        self.ii, self.jj = e(torch.long), e(torch.long)
        self.idx_ii2jj, self.idx_jj2ii = e(torch.long), e(torch.long)
        self.valid_match_j, self.valid_match_i = e(torch.bool), e(torch.bool)
        self.Q_ii2jj, self.Q_jj2ii = e(torch.float32), e(torch.float32)
        self.window_size = self.cfg["window_size"]

    def add_edges(self, ii, jj, idx_ii2jj, idx_jj2ii, valid_match_j, valid_match_i, Q_ii2jj, Q_jj2ii):
        """What add_factors appends after matching and gating (global_opt.py:88-99)."""
        cat = lambda a, b: torch.cat((a.view((0,) + b.shape[1:]) if a.numel() == 0 else a, b), 0)
        self.ii, self.jj = cat(self.ii, ii), cat(self.jj, jj)
        self.idx_ii2jj, self.idx_jj2ii = cat(self.idx_ii2jj, idx_ii2jj), cat(self.idx_jj2ii, idx_jj2ii)
        self.valid_match_j = cat(self.valid_match_j, valid_match_j)
        self.valid_match_i = cat(self.valid_match_i, valid_match_i)
        self.Q_ii2jj, self.Q_jj2ii = cat(self.Q_ii2jj, Q_ii2jj), cat(self.Q_jj2ii, Q_jj2ii)

    def add_factors(self, ii, jj, min_match_frac, is_reloc=False, matcher=None):
        """global_opt.py:30-99.  `matcher(model, frames, ii, jj)` must return what the reference's
        mast3r_match_symmetric returns: (idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj,
        Qji, Qij) with shapes (E,n) i64 / (E,n,1) bool / (E,n,1) f32."""
        matcher = matcher or self.model
        (idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij) = matcher(self.model, self.frames, ii, jj)
        Q_conf = self.cfg["Q_conf"]
        Qj, cnt_j = _be.edge_build(idx_i2j.contiguous(), valid_match_j.contiguous(), Qii.contiguous(),
                                   Qji.contiguous(), Q_conf)
        Qi, cnt_i = _be.edge_build(idx_j2i.contiguous(), valid_match_i.contiguous(), Qjj.contiguous(),
                                   Qij.contiguous(), Q_conf)
        n = idx_i2j.shape[1]
        match_frac_j = cnt_j / n  # valid_j.sum(dim=(1,2)) / nj
        match_frac_i = cnt_i / n
        ii_tensor = torch.as_tensor(ii, device=self.device)
        jj_tensor = torch.as_tensor(jj, device=self.device)
        # both directions must clear the threshold; consecutive keyframes are always kept (:74-77)
        invalid_edges = torch.minimum(match_frac_j, match_frac_i) < min_match_frac
        invalid_edges = (~(ii_tensor == (jj_tensor - 1))) & invalid_edges
        if invalid_edges.any() and is_reloc:
            return False
        valid_edges = ~invalid_edges
        self.add_edges(ii_tensor[valid_edges], jj_tensor[valid_edges], idx_i2j[valid_edges],
                       idx_j2i[valid_edges], valid_match_j[valid_edges], valid_match_i[valid_edges],
                       Qj[valid_edges], Qi[valid_edges])
        return valid_edges.sum() > 0

    def get_unique_kf_idx(self):
        return torch.unique(torch.cat((self.ii, self.jj)), sorted=True)

    def prep_two_way_edges(self):
        return (torch.cat((self.ii, self.jj)), torch.cat((self.jj, self.ii)),
                torch.cat((self.idx_ii2jj, self.idx_jj2ii)),
                torch.cat((self.valid_match_j, self.valid_match_i)),
                torch.cat((self.Q_ii2jj, self.Q_jj2ii)))

    def get_poses_points(self, unique_kf_idx):
        kfs = [self.frames[int(i)] for i in unique_kf_idx]
        Xs = torch.stack([kf.X_canon for kf in kfs])
        T_WCs = Sim3(torch.stack([kf.T_WC.data for kf in kfs]))
        Cs = torch.stack([kf.get_average_conf() for kf in kfs])
        return Xs, T_WCs, Cs

    def _solve(self, calib):
        cfg = self.cfg
        pin = cfg["pin"]
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= pin:
            return None
        Xs, T_WCs, Cs = self.get_poses_points(unique_kf_idx)
        ii, jj, idx_ii2jj, valid_match, Q = self.prep_two_way_edges()
        pose_data = T_WCs.data[:, 0, :]  # contiguous VIEW: the extension updates it in place
        if calib:
            h, w = self.frames[0].img.shape[-2:]
            Xs = constrain_points_to_ray((h, w), Xs, self.K)
            out = _be.gauss_newton_calib(pose_data, Xs, Cs, self.K, ii, jj, idx_ii2jj, valid_match, Q,
                                         h, w, cfg["pixel_border"], cfg["depth_eps"], cfg["sigma_pixel"],
                                         cfg["sigma_depth"], cfg["C_conf"], cfg["Q_conf"],
                                         cfg["max_iters"], cfg["delta_norm"])
        else:
            out = _be.gauss_newton_rays(pose_data, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q,
                                        cfg["sigma_ray"], cfg["sigma_dist"], cfg["C_conf"],
                                        cfg["Q_conf"], cfg["max_iters"], cfg["delta_norm"])
        self.frames.update_T_WCs(T_WCs[pin:], unique_kf_idx[pin:])
        return out

    def solve_GN_rays(self):
        self._solve(calib=False)

    def solve_GN_calib(self):
        self._solve(calib=True)
