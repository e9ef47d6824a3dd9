"""FactorGraph.solve_GN_rays / solve_GN_calib with the reference's behaviour (reference
mast3r_slam/global_opt.py:101-213) on top of mast3r_slam_backends.

Storage.  The reference keeps eight tensors — ii, jj, idx_ii2jj, idx_jj2ii,
valid_match_j, valid_match_i, Q_ii2jj, Q_jj2ii — re-`cat`s all of them on every add_factors
(global_opt.py:88-97: O(E^2) bytes over a run) and concatenates the two directions again before
every solve (prep_two_way_edges, :104-110: ~0.5 GB copied at E = 100).  Here the graph owns ONE
preallocated, capacity-doubling store: `ij (2,cap) i64` and `pk (2,cap,n)` 8-byte packed records
(include/m3r_b200.h m3r_packed_match: idx i32 + Q with valid_match folded into its sign), block 0
the forward direction, block 1 the reverse.  add_factors builds its records IN PLACE in the tail
rows (m3r_edge_build_packed), and the solve hands the two blocks to m3r_gauss_newton_packed as
they lie: no cat, no per-solve copy, 40 instead of 45 B per (edge, point) in the GN kernel.

The reference's attribute names stay readable (`ii`, `jj` are views; `idx_ii2jj`, `valid_match_j`,
`Q_ii2jj`, ... and prep_two_way_edges() unpack on demand for callers that want the reference
layout — the solve never does).

Edge construction (`add_factors`, :30-99): the MASt3R network + symmetric matching is caller-side
(`matcher`, the reference's mast3r_match_symmetric); everything after it — Q = sqrt(Q_src[idx] * Q_dst),
the Q_conf gate, match fractions, edge acceptance and appending — is here, the per-point part as
one CUDA launch per direction.  `add_edges` appends already-built reference-layout tensors.

`frames` is any keyframe store with the reference's protocol: `frames[i]` has `.X_canon (n,3)`,
`.T_WC` (Sim3 with data (1,8)), `.get_average_conf() (n,1)`, `.img (3,h,w)`; and
`frames.update_T_WCs(T_WCs, idx)` (frame.py:309-317)."""
import torch

import mast3r_slam_backends as _be
from mast3r_slam.config import config
from mast3r_slam.geometry import constrain_points_to_ray
from mast3r_slam.sim3 import Sim3


class _EdgeStore:
    """ij (2,cap) i64 and pk (2,cap,n) packed records; rows [0,E) of each block are live."""

    def __init__(self, device):
        self.device = device
        self.E, self.cap, self.n = 0, 0, None
        self.ij = torch.empty((2, 0), dtype=torch.int64, device=device)
        self.pk = None
        self.reallocations = 0  # diagnostics: how often the store had to grow (amortised O(1) per edge)

    def reserve(self, extra, n):
        """Room for `extra` more edges of n points; returns nothing, may move the store once."""
        if self.n is None:
            self.n = int(n)
        elif self.n != int(n):
            raise RuntimeError(f"FactorGraph: edges have {n} points, the graph holds edges of {self.n}")
        need = self.E + int(extra)
        if need <= self.cap:
            return
        cap = max(need, 2 * self.cap, 8)
        ij = torch.empty((2, cap), dtype=torch.int64, device=self.device)
        pk = torch.empty((2, cap, self.n), dtype=torch.int64, device=self.device)
        if self.E:
            ij[:, :self.E] = self.ij[:, :self.E]
            pk[:, :self.E] = self.pk[:, :self.E]
        self.ij, self.pk, self.cap = ij, pk, cap
        self.reallocations += 1

    def tail(self, block, count):
        """The `count` rows after the live ones of a block — where new edges are built in place."""
        return self.pk[block, self.E:self.E + count]

    def commit(self, ii, jj, keep=None):
        """Make the tail rows live.  keep: bool mask over them (None = all); rejected rows are
        squeezed out (a copy of the ACCEPTED rows of this batch only)."""
        cnt = ii.shape[0]
        if keep is not None:
            rows = torch.nonzero(keep, as_tuple=False).flatten()
            k = int(rows.numel())
            if k < cnt:
                for b in range(2):
                    self.pk[b, self.E:self.E + k] = self.pk[b, self.E:self.E + cnt].index_select(0, rows)
                ii, jj, cnt = ii[rows], jj[rows], k
        self.ij[0, self.E:self.E + cnt] = ii
        self.ij[1, self.E:self.E + cnt] = jj
        self.E += cnt
        return cnt


class FactorGraph:
    def __init__(self, model=None, frames=None, K=None, device="cuda"):
        self.model = model
        self.frames = frames
        self.device = device
        self.cfg = config["local_opt"]
        self.K = K
        self._store = _EdgeStore(device)
        self.window_size = self.cfg["window_size"]

    # ------------------------------------------------------------------ the reference's attributes
    @property
    def ii(self):
        return self._store.ij[0, :self._store.E]

    @property
    def jj(self):
        return self._store.ij[1, :self._store.E]

    def _unpacked(self, block):
        st = self._store
        if st.E == 0:
            e = lambda dt: torch.as_tensor([], dtype=dt, device=self.device)
            return e(torch.long), e(torch.bool), e(torch.float32)
        return _be.edge_unpack(st.pk[block, :st.E])

    idx_ii2jj = property(lambda self: self._unpacked(0)[0])
    idx_jj2ii = property(lambda self: self._unpacked(1)[0])
    valid_match_j = property(lambda self: self._unpacked(0)[1])
    valid_match_i = property(lambda self: self._unpacked(1)[1])
    Q_ii2jj = property(lambda self: self._unpacked(0)[2])
    Q_jj2ii = property(lambda self: self._unpacked(1)[2])

    def packed_edges(self):
        """-> (ii (E), jj (E), forward (E,n), reverse (E,n)): VIEWS of the store, what the solve reads."""
        st = self._store
        return self.ii, self.jj, st.pk[0, :st.E], st.pk[1, :st.E]

    # ------------------------------------------------------------------ building the graph
    def add_edges(self, ii, jj, idx_ii2jj, idx_jj2ii, valid_match_j, valid_match_i, Q_ii2jj, Q_jj2ii):
        """What add_factors appends after matching and gating (global_opt.py:88-99), given in the
        reference's layout; packed straight into the store's tail rows."""
        cnt = int(ii.shape[0])
        if cnt == 0:
            return
        st = self._store
        st.reserve(cnt, idx_ii2jj.shape[1])
        _be.edge_pack(idx_ii2jj.contiguous(), valid_match_j.contiguous(), Q_ii2jj.contiguous(), out=st.tail(0, cnt))
        _be.edge_pack(idx_jj2ii.contiguous(), valid_match_i.contiguous(), Q_jj2ii.contiguous(), out=st.tail(1, cnt))
        st.commit(torch.as_tensor(ii, device=self.device, dtype=torch.int64),
                  torch.as_tensor(jj, device=self.device, dtype=torch.int64))

    def _match(self, ii, jj):
        """The reference's own matcher call (global_opt.py:31-53); needs its mast3r_utils (the MASt3R
        network is outside this package's scope)."""
        try:
            from mast3r_slam.mast3r_utils import mast3r_match_symmetric
        except ImportError as e:
            raise RuntimeError(
                "FactorGraph.add_factors: no `matcher` given and mast3r_slam.mast3r_utils (the reference's "
                "MASt3R front end) is not importable; pass matcher=callable(model, frames, ii, jj) returning "
                "mast3r_match_symmetric's 8 outputs") from e
        kf_ii = [self.frames[idx] for idx in ii]
        kf_jj = [self.frames[idx] for idx in jj]
        feat_i = torch.cat([kf.feat for kf in kf_ii]); feat_j = torch.cat([kf.feat for kf in kf_jj])
        pos_i = torch.cat([kf.pos for kf in kf_ii]); pos_j = torch.cat([kf.pos for kf in kf_jj])
        shape_i = [kf.img_true_shape for kf in kf_ii]; shape_j = [kf.img_true_shape for kf in kf_jj]
        return mast3r_match_symmetric(self.model, feat_i, pos_i, feat_j, pos_j, shape_i, shape_j)

    def add_factors(self, ii, jj, min_match_frac, is_reloc=False, matcher=None):
        """global_opt.py:30-99.  `matcher(model, frames, ii, jj)` must return what the reference's
        mast3r_match_symmetric returns: (idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj,
        Qji, Qij) with shapes (E,n) i64 / (E,n,1) bool / (E,n,1) f32; without one the reference's
        own call is made (needs mast3r_slam.mast3r_utils)."""
        if matcher is not None:
            out = matcher(self.model, self.frames, ii, jj)
        else:
            out = self._match(ii, jj)
        (idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij) = out
        Q_conf = self.cfg["Q_conf"]
        cand, n = idx_i2j.shape
        st = self._store
        st.reserve(cand, n)
        # records of BOTH directions are built in place in the store's tail rows
        _, cnt_j = _be.edge_build_packed(idx_i2j.contiguous(), valid_match_j.contiguous(), Qii.contiguous(),
                                         Qji.contiguous(), Q_conf, out=st.tail(0, cand))
        _, cnt_i = _be.edge_build_packed(idx_j2i.contiguous(), valid_match_i.contiguous(), Qjj.contiguous(),
                                         Qij.contiguous(), Q_conf, out=st.tail(1, cand))
        match_frac_j = cnt_j / n  # valid_j.sum(dim=(1,2)) / nj
        match_frac_i = cnt_i / n
        ii_tensor = torch.as_tensor(ii, device=self.device, dtype=torch.int64)
        jj_tensor = torch.as_tensor(jj, device=self.device, dtype=torch.int64)
        # both directions must clear the threshold; consecutive keyframes are always kept (:74-77)
        invalid_edges = torch.minimum(match_frac_j, match_frac_i) < min_match_frac
        invalid_edges = (~(ii_tensor == (jj_tensor - 1))) & invalid_edges
        any_invalid = bool(invalid_edges.any())  # the reference syncs here too (:79, :82)
        if any_invalid and is_reloc:
            return False  # nothing committed: the tail rows stay dead
        added = st.commit(ii_tensor, jj_tensor, keep=(~invalid_edges) if any_invalid else None)
        return added > 0

    # ------------------------------------------------------------------ solving
    def get_unique_kf_idx(self):
        return torch.unique(torch.cat((self.ii, self.jj)), sorted=True)

    def prep_two_way_edges(self):
        """The reference's five tensors (global_opt.py:104-110), unpacked and concatenated ON
        REQUEST for callers that want that layout.  solve_GN_* does not call this."""
        idx_f, vm_f, Q_f = self._unpacked(0)
        idx_r, vm_r, Q_r = self._unpacked(1)
        return (torch.cat((self.ii, self.jj)), torch.cat((self.jj, self.ii)), torch.cat((idx_f, idx_r)),
                torch.cat((vm_f, vm_r)), torch.cat((Q_f, Q_r)))

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
        ii, jj, fwd, rev = self.packed_edges()
        pose_data = T_WCs.data[:, 0, :]  # contiguous VIEW: the extension updates it in place
        if calib:
            h, w = self.frames[0].img.shape[-2:]
            Xs = constrain_points_to_ray((h, w), Xs, self.K)
            prm = _be._params(_be.GN_CALIB, cfg["sigma_pixel"], cfg["sigma_depth"], cfg["C_conf"], cfg["Q_conf"],
                              self.K, h, w, cfg["pixel_border"], cfg["depth_eps"], device=pose_data.device)
        else:
            prm = _be._params(_be.GN_RAYS, cfg["sigma_ray"], cfg["sigma_dist"], cfg["C_conf"], cfg["Q_conf"])
        out = _be.gauss_newton_packed(prm, pose_data, Xs, Cs, ii, jj, fwd, rev, cfg["max_iters"],
                                      cfg["delta_norm"], two_way=True)
        self.frames.update_T_WCs(T_WCs[pin:], unique_kf_idx[pin:])
        return out

    def solve_GN_rays(self):
        self._solve(calib=False)

    def solve_GN_calib(self):
        self._solve(calib=True)
