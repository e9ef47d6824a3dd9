"""Generate tests/golden/tracker_ref.npz by running the reference's OWN tracker code.

Runs in the build container only (needs /root/reference; the GPU box has no copy).  The
reference's `mast3r_slam/tracker.py` is imported unmodified; its un-vendored dependencies are
replaced by stand-ins: `lietorch` -> oracle.tracker.Sim3, and the two modules that
pull in the MASt3R network (`mast3r_slam.frame`, `mast3r_slam.mast3r_utils`) -> empty stubs
(the method under test, FrameTracker.opt_pose_ray_dist_sim3, touches neither).
Config values come from the reference's config/base.yaml through its own loader.

    python tests/golden/gen_tracker_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
sys.path.insert(0, ROOT)

from oracle import tracker as otrk  # noqa: E402
from tests import synth  # noqa: E402


def import_reference_tracker():
    liet = types.ModuleType("lietorch")
    liet.Sim3 = otrk.Sim3
    sys.modules["lietorch"] = liet
    frame = types.ModuleType("mast3r_slam.frame")
    frame.Frame = object
    frame.SharedKeyframes = object
    utils = types.ModuleType("mast3r_slam.mast3r_utils")
    utils.mast3r_match_asymmetric = None
    utils.mast3r_match_symmetric = None
    sys.path.insert(0, REF)
    import mast3r_slam  # noqa: F401  (namespace package of the reference)

    sys.modules["mast3r_slam.frame"] = frame
    sys.modules["mast3r_slam.mast3r_utils"] = utils
    from mast3r_slam.config import config, load_config

    load_config(os.path.join(REF, "config", "base.yaml"))
    from mast3r_slam import tracker as ref_tracker

    return ref_tracker, config


def main():
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    ref_tracker, config = import_reference_tracker()
    trk = ref_tracker.FrameTracker(None, None, "cpu")
    out = {}
    for name, (H, W, seed) in {"a": (48, 64, 1234), "b": (96, 128, 77)}.items():
        d = synth.make_tracker_inputs(H, W, seed=seed, device="cpu")
        T_WCk = otrk.Sim3(torch.tensor([[0.3, -0.1, 0.2, 0.05, 0.02, -0.03, 0.0, 1.1]]))
        T_WCk.data[0, 3:7] /= 1.0  # filled below
        q = torch.tensor([0.05, 0.02, -0.03, 1.0])
        T_WCk.data[0, 3:7] = q / q.norm()
        T_WCf = otrk.Sim3(T_WCk.data.clone())  # start: frame at the keyframe pose => T_CkCf = I
        # capture the iterates the reference code produces through its own solve()
        hist = []
        orig_solve = trk.solve

        def spy(sqrt_info, r, J):
            tau, cost = orig_solve(sqrt_info, r, J)
            hist.append((cost, tau.clone()))
            return tau, cost

        trk.solve = spy
        T_WCf_new, T_CkCf = trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], T_WCf, T_WCk, d["Qk"],
                                                       d["valid"])
        trk.solve = orig_solve
        out[f"{name}_HW"] = np.array([H, W, seed])
        out[f"{name}_T_WCk"] = T_WCk.data.numpy()
        out[f"{name}_T_WCf_out"] = T_WCf_new.data.numpy()
        out[f"{name}_T_CkCf_out"] = T_CkCf.data.numpy()
        out[f"{name}_costs"] = np.array([h[0] for h in hist], np.float64)
        out[f"{name}_taus"] = torch.cat([h[1] for h in hist]).numpy()
        out[f"{name}_T_gt"] = d["T_gt"].numpy()
        print(name, "iterations", len(hist), "T_CkCf", T_CkCf.data.numpy().round(5))
    # calibrated twin: the reference's opt_pose_calib_sim3 (tracker.py:216-266) with the measurement
    # construction of get_points_poses (:143-152) done by the reference's own geometry helpers
    from mast3r_slam import geometry as ref_geom

    for name, (H, W, seed) in {"ca": (48, 64, 1234), "cb": (96, 128, 77)}.items():
        d = synth.make_tracker_calib_inputs(H, W, seed=seed, device="cpu")
        q = torch.tensor([0.05, 0.02, -0.03, 1.0])
        T_WCk = otrk.Sim3(torch.cat([torch.tensor([0.3, -0.1, 0.2]), q / q.norm(), torch.tensor([1.1])])[None])
        T_WCf = otrk.Sim3(T_WCk.data.clone())
        uv_k = ref_geom.get_pixel_coords(1, (H, W), device="cpu", dtype=torch.float32).view(-1, 2)
        meas_k = torch.cat((uv_k, torch.log(d["Xk"][..., 2:3])), dim=-1)
        valid_meas_k = d["Xk"][..., 2:3] > config["tracking"]["depth_eps"]
        meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
        hist = []
        orig_solve = trk.solve

        def spy(sqrt_info, r, J):
            tau, cost = orig_solve(sqrt_info, r, J)
            hist.append((cost, tau.clone()))
            return tau, cost

        trk.solve = spy
        T_WCf_new, T_CkCf = trk.opt_pose_calib_sim3(d["Xf"], d["Xk"], T_WCf, T_WCk, d["Qk"], d["valid"],
                                                     meas_k, valid_meas_k, d["K"], (H, W))
        trk.solve = orig_solve
        out[f"{name}_HW"] = np.array([H, W, seed])
        out[f"{name}_T_WCk"] = T_WCk.data.numpy()
        out[f"{name}_T_WCf_out"] = T_WCf_new.data.numpy()
        out[f"{name}_T_CkCf_out"] = T_CkCf.data.numpy()
        out[f"{name}_costs"] = np.array([h[0] for h in hist], np.float64)
        out[f"{name}_taus"] = torch.cat([h[1] for h in hist]).numpy()
        print(name, "iterations", len(hist), "T_CkCf", T_CkCf.data.numpy().round(5))
    out["cfg_calib"] = np.array([config["tracking"][k] for k in
                                 ("sigma_pixel", "sigma_depth", "pixel_border", "depth_eps")], np.float64)
    out["cfg"] = np.array([config["tracking"][k] for k in
                           ("sigma_ray", "sigma_dist", "huber", "max_iters", "rel_error", "delta_norm")],
                          np.float64)
    path = os.path.join(ROOT, "tests", "golden", "tracker_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main()
