"""FrameTracker.opt_pose_ray_dist_sim3 with the reference's signature (reference
mast3r_slam/tracker.py:173-214).  The reference loops in Python over ~25 PyTorch kernels and two
host syncs per iteration; here the whole Gauss-Newton loop — residuals, Huber weights, 7x7
normal equations, Cholesky, Sim3 retraction, the convergence test — is ONE cooperative CUDA
launch (m3r_track_ray_dist) and one 16-byte status read."""
import torch

import mast3r_slam_backends as _be
from mast3r_slam.config import config


class FrameTracker:
    def __init__(self, model=None, frames=None, device="cuda"):
        self.cfg = config["tracking"]
        self.model = model
        self.keyframes = frames
        self.device = device
        self.idx_f2k = None
        self.last_status = None

    def reset_idx_f2k(self):
        self.idx_f2k = None

    def opt_pose_ray_dist_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid):
        """Xf, Xk (n,3) f32; T_WCf, T_WCk Sim3 with data (1,8); Qk (n,1) f32; valid (n,1) bool
        -> (T_WCf, T_CkCf).  Raises RuntimeError when the 7x7 system is not positive definite,
        like torch.linalg.cholesky in the reference's solve (tracker.py:168; caught at :91)."""
        cfg = self.cfg
        T_CkCf = T_WCk.inv() * T_WCf
        rel = T_CkCf.data.reshape(8).to(torch.float32).contiguous().clone()
        status = _be.track_ray_dist(Xf.contiguous(), Xk.contiguous(), Qk.contiguous(),
                                    valid.contiguous(), rel, cfg["sigma_ray"], cfg["sigma_dist"],
                                    cfg["huber"], cfg["max_iters"], cfg["rel_error"], cfg["delta_norm"])
        iters, converged, failed, _ = status.tolist()  # the solve's only host sync
        self.last_status = dict(iterations=iters, converged=bool(converged))
        if failed:
            raise RuntimeError("opt_pose_ray_dist_sim3: normal equations not positive definite")
        if not converged and iters == cfg["max_iters"]:
            print("max iters reached 0")
        T_CkCf = type(T_WCk)(rel.view(T_WCk.data.shape))
        return T_WCk * T_CkCf, T_CkCf

    def opt_pose_calib_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size):
        """Calibrated twin (tracker.py:216-266): pixel + log-depth residuals against meas_k.
        Same contract as opt_pose_ray_dist_sim3; Xk is unused by the solve, as in the reference."""
        cfg = self.cfg
        T_CkCf = T_WCk.inv() * T_WCf
        rel = T_CkCf.data.reshape(8).to(torch.float32).contiguous().clone()
        status = _be.track_calib(Xf.contiguous(), meas_k.contiguous(), Qk.contiguous(), valid.contiguous(),
                                 valid_meas_k.contiguous(), K, img_size, rel, cfg["sigma_pixel"],
                                 cfg["sigma_depth"], cfg["pixel_border"], cfg["depth_eps"], cfg["huber"],
                                 cfg["max_iters"], cfg["rel_error"], cfg["delta_norm"])
        iters, converged, failed, _ = status.tolist()
        self.last_status = dict(iterations=iters, converged=bool(converged))
        if failed:
            raise RuntimeError("opt_pose_calib_sim3: normal equations not positive definite")
        if not converged and iters == cfg["max_iters"]:
            print("max iters reached 0")
        T_CkCf = type(T_WCk)(rel.view(T_WCk.data.shape))
        return T_WCk * T_CkCf, T_CkCf
