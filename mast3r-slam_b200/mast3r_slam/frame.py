"""Frame with the reference's pointmap bookkeeping (reference mast3r_slam/frame.py:16-108):
`update_pointmap(X, C)` under config["tracking"]["filtering_mode"], `get_average_conf()`.
The per-point modes (weighted_pointmap — the default —, weighted_spherical, indep_conf) run as one
in-place CUDA launch (m3r_pointmap_update); `update_pointmap(X, C, T=...)` additionally fuses the
`T_CkCf.act(Xkf)` that the tracker applies just before (tracker.py:97-98)."""
import dataclasses
from typing import Optional

import torch

import mast3r_slam_backends as _be
from mast3r_slam.config import config
from mast3r_slam.sim3 import Sim3


@dataclasses.dataclass
class Frame:
    frame_id: int
    img: Optional[torch.Tensor] = None
    img_shape: Optional[torch.Tensor] = None
    img_true_shape: Optional[torch.Tensor] = None
    uimg: Optional[torch.Tensor] = None
    T_WC: Optional[Sim3] = None
    X_canon: Optional[torch.Tensor] = None
    C: Optional[torch.Tensor] = None
    feat: Optional[torch.Tensor] = None
    pos: Optional[torch.Tensor] = None
    N: int = 0
    N_updates: int = 0
    K: Optional[torch.Tensor] = None
    score: Optional[torch.Tensor] = None

    def get_score(self, C):  # frame.py:31-39
        if config["tracking"].get("filtering_score", "median") == "median":
            return torch.median(C)
        return torch.mean(C)

    def _replace(self, X, C, T):
        self.X_canon = X.clone() if T is None else T.act(X)
        self.C = C.clone()
        self.N = 1

    def update_pointmap(self, X, C, T=None):
        """X (n,3), C (n,1) f32 CUDA.  T: optional Sim3 applied to X first."""
        mode = config["tracking"].get("filtering_mode", "weighted_pointmap")
        if self.N == 0:
            self._replace(X, C, T)
            self.N_updates = 1
            if mode == "best_score":
                self.score = self.get_score(C)
            return
        if mode == "first":
            if self.N_updates == 1:
                self._replace(X, C, T)
        elif mode == "recent":
            self._replace(X, C, T)
        elif mode == "best_score":
            new_score = self.get_score(C)
            if new_score > self.score:
                self._replace(X, C, T)
                self.score = new_score
        elif mode in _be.POINTMAP_MODES:
            if not (self.X_canon.is_contiguous() and self.C.is_contiguous()):
                self.X_canon, self.C = self.X_canon.contiguous(), self.C.contiguous()
            Tdata = None if T is None else T.data.reshape(8).to(torch.float32).contiguous()
            _be.pointmap_update(self.X_canon, self.C, X.contiguous(), C.contiguous(), mode, Tdata)
            self.N = 1 if mode == "indep_conf" else self.N + 1
        else:
            raise ValueError(f"unknown filtering_mode {mode}")
        self.N_updates += 1

    def get_average_conf(self):
        return self.C / self.N if self.C is not None else None
