"""Generate tests/golden/frontend_ref.npz by running the reference's OWN code:
  * FactorGraph.add_factors (mast3r_slam/global_opt.py:30-99) with the network-side call
    `mast3r_match_symmetric` replaced by a stand-in that returns seeded synthetic match outputs
    (tests/synth.make_symmetric_match_outputs) — everything after the matcher is reference code;
  * Frame.update_pointmap (mast3r_slam/frame.py:41-105) for the per-point filtering modes.
Build container only (needs /root/reference).  Stand-ins for un-vendored dependencies: lietorch ->
oracle.tracker.Sim3, mast3r_slam.mast3r_utils / mast3r_slam_backends -> stubs (not reached).

    python tests/golden/gen_frontend_golden.py
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

H, W, E = 48, 64, 5
II, JJ = [0, 1, 0, 2, 1], [1, 2, 3, 5, 4]   # edges 0,1 consecutive; 2 and 4 weak (4 is rejected, 2 too)
WEAK = (0, 2, 4)                            # edge 0 is weak but consecutive => kept


def import_reference():
    liet = types.ModuleType("lietorch")
    liet.Sim3 = otrk.Sim3
    sys.modules["lietorch"] = liet
    utils = types.ModuleType("mast3r_slam.mast3r_utils")
    utils.mast3r_match_symmetric = None
    utils.mast3r_match_asymmetric = None
    utils.resize_img = None
    sys.modules["mast3r_slam_backends"] = types.ModuleType("mast3r_slam_backends")
    sys.path.insert(0, REF)
    import mast3r_slam  # noqa: F401

    sys.modules["mast3r_slam.mast3r_utils"] = utils
    from mast3r_slam.config import config, load_config

    load_config(os.path.join(REF, "config", "base.yaml"))
    from mast3r_slam import frame as ref_frame
    from mast3r_slam import global_opt as ref_go

    return ref_frame, ref_go, config


class _KF:
    feat = torch.zeros(1, 1)
    pos = torch.zeros(1, 1)
    img_true_shape = torch.tensor([[H, W]])


def main():
    torch.manual_seed(0)
    ref_frame, ref_go, config = import_reference()
    out = {}
    # ---- add_factors
    m = synth.make_symmetric_match_outputs(E, H, W, seed=77, device="cpu", weak=WEAK)

    def fake_matcher(model, feat_i, pos_i, feat_j, pos_j, shape_i, shape_j):
        return (m["idx_i2j"], m["idx_j2i"], m["valid_i2j"], m["valid_j2i"], m["Qii"], m["Qjj"], m["Qji"], m["Qij"])

    ref_go.mast3r_match_symmetric = fake_matcher
    frames = {i: _KF() for i in range(8)}
    for min_frac, tag in ((0.1, "a"), (0.0, "b")):
        g = ref_go.FactorGraph(None, frames, None, "cpu")
        added = g.add_factors(II, JJ, min_frac, is_reloc=False)
        out[f"af_{tag}_added"] = np.array(bool(added))
        for k in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
            out[f"af_{tag}_{k}"] = getattr(g, k).numpy()
        print("add_factors", tag, "kept edges", g.ii.tolist(), g.jj.tolist())
    g = ref_go.FactorGraph(None, frames, None, "cpu")
    out["af_reloc_result"] = np.array(bool(g.add_factors(II, JJ, 0.1, is_reloc=True)))
    out["af_cfg"] = np.array([config["local_opt"]["Q_conf"], H, W, E, 77], np.float64)
    out["af_edges"] = np.array([II, JJ])
    out["af_weak"] = np.array(WEAK)
    # ---- update_pointmap
    d = synth.make_match_inputs(2, H, W, 8, seed=91, device="cpu")
    X0, X1 = d["X11"][0].reshape(-1, 3), d["X21"][1].reshape(-1, 3)
    gen = torch.Generator().manual_seed(5)
    C0 = 1.0 + torch.exp(torch.randn((H * W, 1), generator=gen))
    C1 = 1.0 + torch.exp(torch.randn((H * W, 1), generator=gen))
    for mode in ("weighted_pointmap", "weighted_spherical", "indep_conf"):
        config["tracking"]["filtering_mode"] = mode
        f = ref_frame.Frame(0, None, None, None, None)
        f.update_pointmap(X0, C0)
        f.update_pointmap(X1, C1)
        f.update_pointmap(X0 * 1.01, C1 * 0.5)
        out[f"pm_{mode}_X"] = f.X_canon.numpy()
        out[f"pm_{mode}_C"] = f.C.numpy()
        out[f"pm_{mode}_N"] = np.array([f.N, f.N_updates])
        out[f"pm_{mode}_avg"] = f.get_average_conf().numpy()
    out["pm_X0"], out["pm_X1"], out["pm_C0"], out["pm_C1"] = X0.numpy(), X1.numpy(), C0.numpy(), C1.numpy()
    path = os.path.join(ROOT, "tests", "golden", "frontend_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
