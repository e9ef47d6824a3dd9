"""GPU parity (through the C ABI) of m3r_edge_build / m3r_pointmap_update and their host mirrors
(FactorGraph.add_factors, Frame.update_pointmap) against the reference's own outputs (golden) and,
at 512x384, against the CPU oracle."""
import os

import numpy as np
import pytest
import torch

from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _np(t):
    return t.detach().cpu().numpy()


def test_add_factors_vs_reference_golden(backends):
    from mast3r_slam.global_opt import FactorGraph

    z = np.load(os.path.join(GOLD, "frontend_ref.npz"))
    Q_conf, H, W, E, seed = z["af_cfg"]
    m = synth.make_symmetric_match_outputs(int(E), int(H), int(W), seed=int(seed), device="cpu",
                                           weak=tuple(int(w) for w in z["af_weak"]))
    m = {k: v.to(DEV) for k, v in m.items()}
    ii, jj = (list(map(int, r)) for r in z["af_edges"])

    def matcher(model, frames, ii_, jj_):
        return (m["idx_i2j"], m["idx_j2i"], m["valid_i2j"], m["valid_j2i"], m["Qii"], m["Qjj"], m["Qji"], m["Qij"])

    for tag, frac in (("a", 0.1), ("b", 0.0)):
        g = FactorGraph(None, None, None, DEV)
        added = g.add_factors(ii, jj, frac, matcher=matcher)
        assert bool(added) == bool(z[f"af_{tag}_added"])
        for k in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
            if k.startswith("Q_"):  # golden = torch CPU sqrt, 1 ulp off on ~1 % (oracle/frontend.py)
                assert np.allclose(_np(getattr(g, k)), z[f"af_{tag}_{k}"], rtol=1.2e-7, atol=0), (tag, k)
            else:
                assert np.array_equal(_np(getattr(g, k)), z[f"af_{tag}_{k}"]), (tag, k)
    g = FactorGraph(None, None, None, DEV)
    assert g.add_factors(ii, jj, 0.1, is_reloc=True, matcher=matcher) is False
    assert g.ii.numel() == 0


def test_edge_build_full_size_vs_oracle(backends):
    from oracle import frontend as ofe

    m = synth.make_symmetric_match_outputs(3, seed=5, device=DEV, weak=(1,))
    Q, cnt = backends.edge_build(m["idx_i2j"], m["valid_i2j"], m["Qii"], m["Qji"], 1.5)
    Qo, co = ofe.edge_build(m["idx_i2j"].cpu(), m["valid_i2j"].cpu(), m["Qii"].cpu(), m["Qji"].cpu(), 1.5)
    assert np.array_equal(_np(Q), _np(Qo)) and np.array_equal(_np(cnt), _np(co))
    # ragged n (not a multiple of the 1024-point CTA tile) and E = 0
    idx = m["idx_i2j"][:, :1001].clamp(max=1000).contiguous()
    Q2, c2 = backends.edge_build(idx, m["valid_i2j"][:, :1001].contiguous(), m["Qii"][:, :1001].contiguous(),
                                 m["Qji"][:, :1001].contiguous(), 1.5)
    Q2o, c2o = ofe.edge_build(idx.cpu(), m["valid_i2j"][:, :1001].cpu(), m["Qii"][:, :1001].cpu(),
                              m["Qji"][:, :1001].cpu(), 1.5)
    assert np.array_equal(_np(Q2), _np(Q2o)) and np.array_equal(_np(c2), _np(c2o))
    Q0, c0 = backends.edge_build(idx[:0], m["valid_i2j"][:0, :1001].contiguous(), m["Qii"][:0, :1001].contiguous(),
                                 m["Qji"][:0, :1001].contiguous(), 1.5)
    assert Q0.shape == (0, 1001, 1) and c0.numel() == 0


def test_update_pointmap_vs_reference_golden(backends):
    from mast3r_slam.config import config, reset_config
    from mast3r_slam.frame import Frame

    z = np.load(os.path.join(GOLD, "frontend_ref.npz"))
    X0, X1, C0, C1 = (torch.from_numpy(z[k]).to(DEV) for k in ("pm_X0", "pm_X1", "pm_C0", "pm_C1"))
    try:
        for mode in ("weighted_pointmap", "weighted_spherical", "indep_conf"):
            config["tracking"]["filtering_mode"] = mode
            f = Frame(0)
            f.update_pointmap(X0, C0)
            f.update_pointmap(X1, C1)
            f.update_pointmap(X0 * 1.01, C1 * 0.5)
            assert [f.N, f.N_updates] == list(z[f"pm_{mode}_N"])
            if mode == "weighted_spherical":  # atan2/acos/sin/cos: CUDA libm vs glibc, few ulp
                assert np.allclose(_np(f.X_canon), z[f"pm_{mode}_X"], rtol=2e-5, atol=2e-6)
            else:                              # same operation order, no FMA contraction: bit-exact
                assert np.array_equal(_np(f.X_canon), z[f"pm_{mode}_X"])
            assert np.array_equal(_np(f.C), z[f"pm_{mode}_C"])
            # C / N is a torch op on both sides (CUDA divides by a scalar via the reciprocal): 1 ulp
            assert np.allclose(_np(f.get_average_conf()), z[f"pm_{mode}_avg"], rtol=1.2e-7, atol=0)
    finally:
        reset_config()


def test_update_pointmap_fused_act_full_size(backends):
    """T.act(X) fused in front of the weighted update (tracker.py:97-98) vs act-then-update."""
    from mast3r_slam.sim3 import Sim3
    from oracle import frontend as ofe

    d = synth.make_tracker_inputs(seed=4, device=DEV)
    n = d["Xk"].shape[0]
    C0 = d["Qk"].clone(); C1 = (d["Qk"] * 0.7 + 0.2).contiguous()
    T = Sim3(d["T_gt"][None].clone())
    Xc, Cc = d["Xk"].clone(), C0.clone()
    backends.pointmap_update(Xc, Cc, d["Xf"], C1, "weighted_pointmap", T.data.reshape(8).contiguous())
    Xo, Co = ofe.update_pointmap(d["Xk"].cpu(), C0.cpu(), T.act(d["Xf"]).cpu(), C1.cpu(), "weighted_pointmap")
    assert np.allclose(_np(Xc), _np(Xo), rtol=1e-6, atol=1e-6) and np.array_equal(_np(Cc), _np(Co))
    # size-independent property: equal weights and equal points are a fixed point
    Xs, Cs = d["Xk"].clone(), C0.clone()
    backends.pointmap_update(Xs, Cs, d["Xk"], C0, "weighted_pointmap")
    assert np.allclose(_np(Xs), _np(d["Xk"]), rtol=2e-7, atol=0) and np.array_equal(_np(Cs), _np(C0 + C0))
    with pytest.raises(RuntimeError):
        backends.pointmap_update(Xs, Cs, d["Xk"][:10].contiguous(), C0, "weighted_pointmap")
    assert n == 384 * 512
