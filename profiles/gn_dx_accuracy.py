"""Where does |dx_ours - dx_reference| of the first GN iterations come from?  (BASELINE config 3/4.)
Compares, for ONE iteration from the same poses: the reference extension, this library, and the oracle's
restatement of the reference's per-point fp32 arithmetic with EXACT (fp64) summation — i.e. what the
reference would return if its sums did not round.  Run on the GPU box:
    python profiles/gn_dx_accuracy.py [out.json]
With M3R_B200_LIB pointing at a -DM3R_GN_PRECISE build the per-point intrinsics are IEEE instead of MUFU."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from oracle import build_ref  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from tests import synth  # noqa: E402

dev = "cuda:0"
ref = build_ref.load_ref()
H, W = synth.H0, synth.W0
out = {"lib": be.LIBRARY_PATH}


def np_(t):
    return t.detach().cpu().numpy()


def solve64(Hs64, gs64, ii_opt, jj_opt, nfree):
    D = 7 * nfree
    A, b = np.zeros((D, D)), np.zeros(D)
    M = ii_opt.shape[0]
    for e in range(M):
        i, j = int(ii_opt[e]), int(jj_opt[e])
        si, sj = slice(7 * i, 7 * i + 7), slice(7 * j, 7 * j + 7)
        if i >= 0:
            A[si, si] += Hs64[0, e]; b[si] += gs64[0, e]
        if j >= 0:
            A[sj, sj] += Hs64[3, e]; b[sj] += gs64[1, e]
        if i >= 0 and j >= 0:
            A[si, sj] += Hs64[1, e]; A[sj, si] += Hs64[2, e]
    return (-np.linalg.solve(A, b)).reshape(nfree, 7), np.linalg.cond(A)


for name, N, counts, calib, seed in (("config3_rays", 32, None, False, 1236),
                                     ("config4_calib", 64, {1: None, 2: 62, 5: 59, "random": 16}, True, 1237)):
    g = synth.make_graph(N=N, counts=counts, seed=seed, device=dev)
    if calib:
        g["Xs"] = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])

    def run(mod, T, iters):
        if calib:
            o = mod.gauss_newton_calib(T, g["Xs"], g["Cs"], g["K"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                       H, W, -10, 1e-6, 1.0, 10.0, 0.0, 1.5, iters, 0.0)
        else:
            o = mod.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
                                      0.003, 10.0, 0.0, 1.5, iters, 0.0)
        torch.cuda.synchronize()
        return o[0]
    rows = []
    T_start = g["Twc"].clone()
    for it in range(3):
        Tr, Tn = T_start.clone(), T_start.clone()
        dx_r, dx_n = np_(run(ref, Tr, 1)), np_(run(be, Tn, 1))
        if calib:
            oprm = orc.make_params(orc.MODE_CALIB, 1.0, 10.0, 0.0, 1.5, np_(g["K"]), H, W, -10, 1e-6)
        else:
            oprm = orc.make_params(orc.MODE_RAYS, 0.003, 10.0, 0.0, 1.5)
        _, ie, je, io, jo = orc.edge_rows(np_(g["ii"]), np_(g["jj"]))
        _, _, Hs64, gs64 = orc.gn_blocks(oprm, np_(T_start), np_(g["Xs"]), np_(g["Cs"]), ie, je, np_(g["idx"]),
                                         np_(g["valid"]), np_(g["Q"]), want64=True)
        dx_x, cond = solve64(Hs64, gs64, io, jo, N - 1)
        rows.append(dict(iteration=it + 1, dx_max=float(np.abs(dx_x).max()), cond=float(cond),
                         ours_vs_ref=float(np.abs(dx_n - dx_r).max()),
                         ours_vs_exact_sum=float(np.abs(dx_n - dx_x).max()),
                         ref_vs_exact_sum=float(np.abs(dx_r - dx_x).max())))
        print(name, rows[-1], flush=True)
        T_start = Tr  # follow the reference's trajectory: every iteration compares from the SAME poses
    out[name] = rows
s = json.dumps(out, indent=1)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(s + "\n")
