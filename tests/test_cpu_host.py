"""CPU tests of the host side: the C-ABI library loads and exports every symbol the header
declares (no compute without a GPU), the product refuses CPU tensors loudly, the host-side
mirror of the reference's caller API (config loader, index helpers, FactorGraph bookkeeping,
Sim3 algebra) behaves like the reference, and the edge-sharding / block-exchange logic of the
multi-GPU solve works over a world_size-2 gloo group."""
import ctypes
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(lib_path, backends):
    hdr = open(os.path.join(ROOT, "include", "m3r_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(m3r_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 20
    lib = ctypes.CDLL(lib_path)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/m3r_b200.h but not exported"
    assert sorted(backends.EXPORTS) == declared  # the Python binding covers the whole header
    dyn = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (m3r_[a-z0-9_]+)", dyn))
    assert exported == set(declared)  # nothing undeclared leaks out of the C ABI
    assert lib.m3r_version() == 100
    lib.m3r_error_string.restype = ctypes.c_char_p
    assert b"workspace" in lib.m3r_error_string(-3)
    # pure host-side queries work without a device
    lib.m3r_gn_workspace_bytes.restype = ctypes.c_size_t
    small = lib.m3r_gn_workspace_bytes(32, 200, 200, 196608)
    assert small > 32 * 196608 * 16
    assert lib.m3r_gn_workspace_bytes(32, 200, 100, 196608) < small
    assert lib.m3r_gn_workspace_bytes(0, 1, 1, 1) == 0


def test_product_has_no_cpu_path(backends):
    with pytest.raises(RuntimeError, match="CUDA"):
        backends.iter_proj(torch.zeros(1, 8, 8, 9), torch.zeros(1, 4, 3), torch.zeros(1, 4, 2), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="CUDA"):
        backends.refine_matches(torch.zeros(1, 8, 8, 24).half(), torch.zeros(1, 4, 24).half(),
                                torch.zeros(1, 4, 2, dtype=torch.long), 3, 5)
    with pytest.raises(RuntimeError, match="CUDA"):
        backends.gauss_newton_rays(torch.zeros(2, 8), torch.zeros(2, 4, 3), torch.zeros(2, 4, 1),
                                   torch.zeros(2, dtype=torch.long), torch.ones(2, dtype=torch.long),
                                   torch.zeros(2, 4, dtype=torch.long), torch.zeros(2, 4, 1, dtype=torch.bool),
                                   torch.zeros(2, 4, 1), 0.003, 10.0, 0.0, 1.5, 10, 1e-8)
    src = open(os.path.join(ROOT, "mast3r-slam_b200", "mast3r_slam_backends", "__init__.py")).read()
    assert "oracle" not in src.replace("gn_blocks_reference_layout", "")  # product never routes to the oracle
    for f in os.listdir(os.path.join(ROOT, "mast3r-slam_b200", "mast3r_slam")):
        if f.endswith(".py"):
            assert "import oracle" not in open(os.path.join(ROOT, "mast3r-slam_b200", "mast3r_slam", f)).read()


def test_config_loader_inherit_and_scientific_floats(backends, tmp_path):
    from mast3r_slam import config as cfgmod

    cfgmod.reset_config()
    assert cfgmod.config["matching"]["lambda_init"] == 1e-8 and cfgmod.config["local_opt"]["max_iters"] == 10
    (tmp_path / "base.yaml").write_text("use_calib: False\nmatching:\n  max_iter: 10\n  lambda_init: 1e-8\n"
                                        "tracking:\n  sigma_dist: 1e+1\n")
    (tmp_path / "child.yaml").write_text(f"inherit: {tmp_path / 'base.yaml'}\nuse_calib: True\n"
                                         "matching:\n  max_iter: 7\n")
    cfgmod.load_config(str(tmp_path / "child.yaml"))
    c = cfgmod.config
    assert c["use_calib"] is True and c["matching"]["max_iter"] == 7
    assert isinstance(c["matching"]["lambda_init"], float) and c["matching"]["lambda_init"] == 1e-8
    assert c["tracking"]["sigma_dist"] == 10.0 and c["tracking"]["huber"] == 1.345  # default kept
    cfgmod.reset_config()
    assert cfgmod.config["matching"]["max_iter"] == 10 and cfgmod.config["use_calib"] is False


def test_index_helpers_and_graph_bookkeeping(backends):
    from mast3r_slam import matching
    from mast3r_slam.global_opt import FactorGraph

    idx = torch.tensor([[0, 5, 513, 196607]])
    p = matching.lin_to_pixel(idx, 512)
    assert p.tolist() == [[[0, 0], [5, 0], [1, 1], [511, 383]]]
    assert torch.equal(matching.pixel_to_lin(p, 512), idx)
    # the capacity-doubling edge store (pure torch bookkeeping; packing itself is a CUDA kernel,
    # tests/test_gpu_frontend.py): records built in the tail rows, rejected rows squeezed out,
    # growth keeps the live rows, no per-append re-cat of the whole graph
    from mast3r_slam.global_opt import _EdgeStore

    st, n = _EdgeStore("cpu"), 6
    st.reserve(3, n)
    assert st.cap >= 3 and st.pk.shape == (2, st.cap, n) and st.E == 0
    for b in range(2):
        st.tail(b, 3).copy_(torch.arange(3 * n).reshape(3, n) + 1000 * b)
    assert st.commit(torch.tensor([4, 9, 7]), torch.tensor([9, 2, 8]), keep=torch.tensor([True, False, True])) == 2
    assert st.E == 2 and st.ij[:, :2].tolist() == [[4, 7], [9, 8]]
    assert st.pk[1, 1].tolist() == list(range(1000 + 2 * n, 1000 + 3 * n))  # row 2 moved up
    base, grew = st.pk.data_ptr(), st.reallocations
    st.reserve(1, n)                                    # fits: the store does not move
    assert st.pk.data_ptr() == base and st.reallocations == grew
    st.reserve(st.cap, n)                               # does not fit: one move, at least doubled
    assert st.reallocations == grew + 1 and st.cap >= 2 * 8 and st.E == 2
    assert st.pk[0, 0].tolist() == list(range(n)) and st.ij[0, :2].tolist() == [4, 7]
    with pytest.raises(RuntimeError):
        st.reserve(1, n + 1)                            # all edges of a graph have the same n
    g = FactorGraph(frames=None, device="cpu")
    assert g.ii.numel() == 0 and g.idx_ii2jj.numel() == 0 and g.get_unique_kf_idx().numel() == 0
    with pytest.raises(RuntimeError, match="CUDA"):     # building edges is CUDA work: no CPU path
        g.add_edges(torch.tensor([4]), torch.tensor([9]), torch.zeros(1, n, dtype=torch.long),
                    torch.zeros(1, n, dtype=torch.long), torch.ones(1, n, 1, dtype=torch.bool),
                    torch.ones(1, n, 1, dtype=torch.bool), torch.ones(1, n, 1), torch.ones(1, n, 1))
    with pytest.raises(RuntimeError, match="matcher"):  # no matcher and no reference front end
        g.add_factors([0], [1], 0.1)
    # fewer keyframes than `pin` => solve is a no-op (global_opt.py:125-127)
    assert FactorGraph(frames=None, device="cpu")._solve(calib=False) is None


def test_sim3_mirror_algebra_matches_oracle_shim(backends):
    from mast3r_slam.sim3 import Sim3
    from oracle import tracker as otrk

    g = torch.Generator().manual_seed(0)
    d = torch.randn(5, 1, 8, generator=g)
    d[..., 3:7] /= d[..., 3:7].norm(dim=-1, keepdim=True)
    d[..., 7] = d[..., 7].abs() + 0.5
    a, b = Sim3(d[:2]), Sim3(d[2:4])
    oa, ob = otrk.Sim3(d[:2]), otrk.Sim3(d[2:4])
    X = torch.randn(2, 7, 3, generator=g)
    assert torch.allclose((a.inv() * b).data, (oa.inv() * ob).data, atol=1e-6)
    assert torch.allclose(a.act(X), oa.act(X), atol=1e-6)
    assert torch.allclose((a * a.inv()).data, Sim3.Identity(2, 1, device="cpu").data, atol=1e-6)
    assert torch.allclose(a.matrix(), oa.matrix(), atol=1e-6)
    assert a[0].data.shape == (1, 8) and a.shape == (2, 1)


def test_constrain_points_to_ray(backends):
    from mast3r_slam.geometry import constrain_points_to_ray
    from tests import synth

    H, W = 6, 8
    K = synth.intrinsics(H, W)
    Xs = torch.rand(2, H * W, 3) + 1.0
    out = constrain_points_to_ray((H, W), Xs, K)
    assert torch.allclose(out, synth.constrain_points_to_ray(Xs, H, W, K), atol=1e-6)
    assert torch.equal(out[..., 2], Xs[..., 2])
    uv = (K @ out[0].T).T
    u = (uv[:, 0] / uv[:, 2]).view(H, W)
    assert torch.allclose(u[3], torch.arange(W, dtype=torch.float32), atol=1e-4)


def test_shard_edges_partition(backends):
    for M, world in [(200, 1), (200, 2), (200, 8), (8192, 8), (10, 4), (3, 8)]:
        spans = [backends.shard_edges(M, world, r) for r in range(world)]
        counts = spans[0][2]
        assert sum(counts) == M and max(counts) - min(counts) <= 1
        cursor = 0
        for r, (b, c, _) in enumerate(spans):
            assert b == cursor and c == counts[r]
            cursor += c


_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["M3R_ROOT"]); sys.path.insert(0, os.path.join(os.environ["M3R_ROOT"], "mast3r-slam_b200"))
import mast3r_slam_backends as be
from oracle import oracle as orc
from tests import synth
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{os.environ['M3R_PORT']}",
                        rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
g = synth.make_graph(N=4, counts={1: None, 2: None, 3: None, "random": 0}, H=12, W=16, seed=3, device="cpu")
M = g["M"]
begin, count, counts = be.shard_edges(M, world, rank)
prm = orc.make_params(0, 0.003, 10.0, 0.0, 1.5)
_, ie, je, io, jo = orc.edge_rows(g["ii"].numpy(), g["jj"].numpy())
sl = slice(begin, begin + count)
# each rank computes ONLY its slice of per-edge blocks (the oracle stands in for the CUDA
# accumulation kernel, which needs a GPU); the exchange + replicated solve is the logic under test
Hs, gs, Hs64, gs64 = orc.gn_blocks(prm, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie[sl], je[sl],
                                   g["idx"].numpy()[sl], g["valid"].numpy()[sl], g["Q"].numpy()[sl], want64=True)
blocks = torch.zeros((M, 56), dtype=torch.float64)
blocks[sl, :49] = torch.from_numpy(Hs64[3].reshape(count, 49))
blocks[sl, 49:] = torch.from_numpy(gs64[1])
be.exchange_blocks(blocks, begin, count, counts)
Hs_all, gs_all = be.gn_blocks_reference_layout(blocks)
dx, ok = orc.sparse_block_solve(Hs_all.numpy().astype(np.float32), gs_all.numpy().astype(np.float32), io, jo, g["N"] - 1)
np.save(os.path.join(os.environ["M3R_OUT"], f"dx_{world}_{rank}.npy"), dx)
np.save(os.path.join(os.environ["M3R_OUT"], f"blocks_{world}_{rank}.npy"), blocks.numpy())
dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_exchange_gloo(backends, orc, tmp_path, world):
    """world_size > 1 over gloo: equal (12 edges / 2 or 3) and unequal slices reproduce the
    single-process blocks and the single-process dx bit-for-bit on every rank."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    outs = {}
    for w in (1, world):
        port = _free_port()
        procs = []
        for r in range(w):
            env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(w), M3R_PORT=str(port), M3R_ROOT=ROOT,
                       M3R_OUT=str(tmp_path), OMP_NUM_THREADS="1")
            procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                          stderr=subprocess.STDOUT, text=True))
        for p in procs:
            out, _ = p.communicate(timeout=300)
            assert p.returncode == 0, out[-2000:]
        outs[w] = [(np.load(tmp_path / f"dx_{w}_{r}.npy"), np.load(tmp_path / f"blocks_{w}_{r}.npy")) for r in range(w)]
    dx1, b1 = outs[1][0]
    assert np.abs(dx1).max() > 0
    for dx, b in outs[world]:
        assert np.array_equal(b, b1) and np.array_equal(dx, dx1)


def test_headless_viewer_host_side(backends):
    """Host logic of the viewer mirror: uniform defaults of the C ABI equal the shader sources'
    (and the oracle's), colour packing, GL matrices.  No device work."""
    import math

    from mast3r_slam import visualization as viz
    from oracle import render as orender

    o, ref = backends.render_opts(), orender.default_opts()
    for name, _ in ref._fields_:
        a, b = getattr(o, name), getattr(ref, name)
        if hasattr(a, "__len__"):
            assert list(a) == list(b), name
        else:
            assert a == b, name
    assert (o.radius, o.slant_threshold, o.spec) == (pytest.approx(0.003), pytest.approx(0.1), 32.0)
    assert backends.render_opts(radius=0.01, phong=(1, 0, 0)).phong[0] == 1.0
    with pytest.raises(TypeError):
        backends.render_opts(no_such_uniform=1)
    assert viz.hex2rgba("#1E2326") == 0xFF26231E  # R in the low byte
    P = viz.perspective(90.0, 2.0, 0.1, 10.0)
    assert P[1, 1] == pytest.approx(1.0) and P[0, 0] == pytest.approx(0.5) and P[3, 2] == -1.0
    # a point on the optical axis at the near / far plane maps to NDC z = -1 / +1
    for z, want in ((-0.1, -1.0), (-10.0, 1.0)):
        c = P @ torch.tensor([0.0, 0.0, z, 1.0])
        assert float(c[2] / c[3]) == pytest.approx(want, abs=1e-5)
    V = viz.look_at((0.0, 0.0, -1.0), (0.0, 0.0, 2.0))
    p = V @ torch.tensor([0.0, 0.0, 2.0, 1.0])          # the target lies 3 units down the -z axis
    assert torch.allclose(p, torch.tensor([0.0, 0.0, -3.0, 1.0]), atol=1e-6)
    up = V @ torch.tensor([0.0, -1.0, 0.0, 0.0])         # the SLAM world's -y is up on screen
    assert float(up[1]) == pytest.approx(1.0)
    # a camera sitting at a pose looks along its +z: view = diag(1,-1,-1,1) * T^-1
    from mast3r_slam.sim3 import Sim3
    T = Sim3(torch.tensor([[0.3, -0.2, 0.5, 0.0, math.sin(0.2), 0.0, math.cos(0.2), 1.7]]))
    Vc = viz.view_from_camera_pose(T)
    fwd = T.matrix().reshape(4, 4).cpu() @ torch.tensor([0.0, 0.0, 1.0, 1.0])  # 1.7 units ahead in the world
    q = Vc @ fwd
    assert torch.allclose(q[:3], torch.tensor([0.0, 0.0, -1.7]), atol=1e-5)
