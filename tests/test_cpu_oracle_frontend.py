"""CPU: the oracle restatement of add_factors' edge tensors and of Frame.update_pointmap against the
golden fixture produced by the reference's own code (tests/golden/gen_frontend_golden.py)."""
import os

import numpy as np
import torch

from oracle import frontend as ofe
from tests import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _inputs(z):
    Q_conf, H, W, E, seed = z["af_cfg"]
    m = synth.make_symmetric_match_outputs(int(E), int(H), int(W), seed=int(seed), device="cpu",
                                           weak=tuple(int(w) for w in z["af_weak"]))
    return m, float(Q_conf), int(H) * int(W)


def test_edge_build_and_gate_vs_reference_add_factors():
    z = np.load(os.path.join(GOLD, "frontend_ref.npz"))
    m, Q_conf, n = _inputs(z)
    ii, jj = z["af_edges"]
    Qj, cj = ofe.edge_build(m["idx_i2j"], m["valid_i2j"], m["Qii"], m["Qji"], Q_conf)
    Qi, ci = ofe.edge_build(m["idx_j2i"], m["valid_j2i"], m["Qjj"], m["Qij"], Q_conf)
    for tag, frac in (("a", 0.1), ("b", 0.0)):
        keep = ofe.add_factors_gate(ii, jj, cj, ci, n, frac)
        assert np.array_equal(np.asarray(ii)[keep.numpy()], z[f"af_{tag}_ii"])
        assert np.array_equal(np.asarray(jj)[keep.numpy()], z[f"af_{tag}_jj"])
        # the golden ran torch's CPU sqrt, which is off by one ulp on ~1 % of inputs (oracle/frontend.py)
        assert np.allclose(Qj[keep].numpy(), z[f"af_{tag}_Q_ii2jj"], rtol=1.2e-7, atol=0)
        assert np.allclose(Qi[keep].numpy(), z[f"af_{tag}_Q_jj2ii"], rtol=1.2e-7, atol=0)
        assert (Qj[keep].numpy() != z[f"af_{tag}_Q_ii2jj"]).mean() < 0.02
        assert np.array_equal(m["idx_i2j"][keep].numpy(), z[f"af_{tag}_idx_ii2jj"])
        assert np.array_equal(m["valid_j2i"][keep].numpy(), z[f"af_{tag}_valid_match_i"])
    assert bool(z["af_a_added"]) and not bool(z["af_reloc_result"])


def test_update_pointmap_vs_reference_frame():
    z = np.load(os.path.join(GOLD, "frontend_ref.npz"))
    X0, X1, C0, C1 = (torch.from_numpy(z[k]) for k in ("pm_X0", "pm_X1", "pm_C0", "pm_C1"))
    for mode in ("weighted_pointmap", "weighted_spherical", "indep_conf"):
        X, C = X0.clone(), C0.clone()
        X, C = ofe.update_pointmap(X, C, X1, C1, mode)
        X, C = ofe.update_pointmap(X, C, X0 * 1.01, C1 * 0.5, mode)
        assert np.array_equal(X.numpy(), z[f"pm_{mode}_X"]), mode
        assert np.array_equal(C.numpy(), z[f"pm_{mode}_C"]), mode
