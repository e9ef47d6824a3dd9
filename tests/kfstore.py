"""A minimal keyframe store with the protocol FactorGraph and FrameTracker expect of the reference's
SharedKeyframes (reference mast3r_slam/frame.py:215-317): preallocated per-keyframe buffers,
`store[i]` -> Frame, `update_T_WCs(T_WCs, idx)` (frame.py:309-311: `self.T_WC[idx] = T_WCs.data`).
Test infrastructure only (the reference's store is shared-memory IPC plumbing, out of scope)."""
import torch

from mast3r_slam.frame import Frame
from mast3r_slam.sim3 import Sim3


class KeyframeStore:
    def __init__(self, X, C, T_WC, img_hw, N=None):
        """X (K,n,3), C (K,n,1) summed confidences, T_WC (K,8), N (K,) update counts (default 1)."""
        K = X.shape[0]
        self.X, self.C = X.contiguous(), C.contiguous()
        self.T_WC = T_WC.clone().view(K, 1, 8)
        self.N = torch.ones(K, dtype=torch.int64) if N is None else N
        self.img = torch.zeros((K, 3) + tuple(img_hw), device=X.device)
        self.updates = 0

    def __len__(self):
        return self.X.shape[0]

    def __getitem__(self, i):
        i = int(i)
        return Frame(frame_id=i, img=self.img[i], T_WC=Sim3(self.T_WC[i]), X_canon=self.X[i], C=self.C[i],
                     N=int(self.N[i]))

    def update_T_WCs(self, T_WCs, idx):
        self.T_WC[idx] = T_WCs.data
        self.updates += 1
