"""m3r_track_ray_dist_noalloc (the entry point with static device scratch, include/m3r_b200.h)
against the existing m3r_track_ray_dist, through the C ABI, on calls made back to back.

Collected last on purpose.  Passes alone in a fresh process —
the first call then performs the per-device cudaGetSymbolAddress set-up — and after the other
tests.  The entry point has no counterpart in the reference."""
import numpy as np
import pytest
import torch

from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CFG = (0.003, 10.0, 1.345, 50, 1e-3, 1e-3)  # sigma_ray, sigma_dist, huber_k, max_iters, rel_error, delta_norm
POSE_TOL = 2e-5  # the tracker tests' pose tolerance (tests/test_gpu_gn.py)


def _identity():
    return torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float32, device=DEV)


def _both(backends, d):
    """-> (T, status, cost) of the workspace entry point and of the no-allocation one."""
    T_ws = _identity()
    status_ws, cost_ws = backends.track_ray_dist(d["Xf"], d["Xk"], d["Qk"], d["valid"], T_ws, *CFG, want_cost=True)
    T_na = _identity()
    status_na = torch.full((4,), -7, dtype=torch.int32, device=DEV)
    cost_na = torch.zeros((CFG[3],), dtype=torch.float32, device=DEV)
    rc = backends.track_ray_dist_noalloc(d["Xf"], d["Xk"], d["Qk"], d["valid"], T_na, status_na, *CFG, cost=cost_na)
    assert rc == 0
    torch.cuda.synchronize()
    return (T_ws, status_ws, cost_ws), (T_na, status_na, cost_na)


def test_noalloc_tracker_equals_workspace_tracker_two_calls_in_a_row(backends):
    # different sizes and contents back to back (the static array keeps the previous call's
    # partials: a second, smaller problem must not see them), then the first problem again
    cases = [synth.make_tracker_inputs(seed=1234, device=DEV),           # 512x384, 148 CTAs
             synth.make_tracker_inputs(48, 64, seed=77, device=DEV),     # 12 CTAs
             synth.make_tracker_inputs(seed=1234, device=DEV)]
    first = None
    for k, d in enumerate(cases):
        (T_ws, st_ws, c_ws), (T_na, st_na, c_na) = _both(backends, d)
        its, conv, fail, _ = st_ws.tolist()
        assert st_na.tolist() == st_ws.tolist() and fail == 0 and conv == 1, (k, st_na.tolist(), st_ws.tolist())
        assert float((T_na - T_ws).abs().max()) <= POSE_TOL, k
        assert np.allclose(c_na.cpu().numpy()[:its], c_ws.cpu().numpy()[:its], rtol=2e-4), k
        # same kernel, same grid, same fixed-order reduction: in fact the same bits
        assert torch.equal(T_na, T_ws), k
        if k == 0:
            first = T_na.clone()
    assert torch.equal(first, T_na)  # call 3 == call 1 although call 2 ran in between


def test_noalloc_tracker_rejects_sizes_beyond_its_maximum(backends):
    d = synth.make_tracker_inputs(48, 64, seed=5, device=DEV)
    T = _identity()
    status = torch.full((4,), -7, dtype=torch.int32, device=DEV)
    launches = backends.launch_count()
    rc = backends.track_ray_dist_noalloc(d["Xf"], d["Xk"], d["Qk"], d["valid"], T, status, *CFG,
                                         n=backends.TRACK_NOALLOC_MAX_N + 1)
    torch.cuda.synchronize()
    assert rc == backends.ETOOLARGE
    assert b"maximum" in backends._lib.m3r_error_string(rc)
    assert backends.launch_count() == launches          # nothing was launched
    assert status.tolist() == [-7] * 4 and torch.equal(T, _identity())
