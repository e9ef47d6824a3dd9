"""Worker of tests/test_gpu_multi.py (run under torchrun, one rank per GPU): the edge-sharded GN solve with
(a) the NCCL all-gather exchange and (b) the exchange fused into the kernels over NVLink peer memory must both
reproduce the single-GPU poses bit for bit.  Prints 'MGPU_OK' on every rank when they do."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for N, H, W, mode in ((9, 48, 64, "rays"), (6, 96, 128, "calib"), (3, 24, 32, "rays")):  # last: fewer edges than ranks x2
    g = synth.make_graph(N=N, counts={1: None, 2: max(N - 3, 1), "random": 1 if N > 4 else 0}, H=H, W=W, seed=77 + N,
                         device=dev)
    if mode == "calib":
        g["Xs"] = synth.constrain_points_to_ray(g["Xs"], H, W, g["K"])
        prm = be._params(be.GN_CALIB, 1.0, 10.0, 0.0, 1.5, g["K"], H, W, -10, 1e-6)
    else:
        prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
    T1 = g["Twc"].clone()
    be._gauss_newton(prm, T1, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], 6, 0.0)
    e0, cnt, counts = be.shard_edges(g["M"], world, rank)
    sl = slice(e0, e0 + cnt)
    loc = [g[k][sl].contiguous() for k in ("idx", "valid", "Q")]
    for p2p in (False, True, True):  # the p2p path twice: buffers and epochs are reused across solves
        T = g["Twc"].clone()
        be.gauss_newton_sharded(prm, T, g["Xs"], g["Cs"], g["ii"], g["jj"], loc[0], loc[1], loc[2], e0, 6, 0.0,
                                edge_counts=counts, p2p=p2p)
        torch.cuda.synchronize()
        same = torch.equal(T, T1)
        moved = not torch.equal(T, g["Twc"])
        if not (same and moved):
            print(f"rank {rank}: N={N} {mode} p2p={p2p} same={same} moved={moved} "
                  f"maxdiff={float((T - T1).abs().max()):.3e}", flush=True)
            ok = False
    # skewed ranks: in every iteration a different rank reaches its accumulation ~300 us late and another
    # one its solve step, so fast ranks push iteration k+1 into the peers' buffers while slow ones still
    # assemble iteration k (the two-buffer / epoch protocol of PeerBlocks must absorb that)
    peers = be.PeerBlocks.get(g["M"], dev)
    for pattern in range(2):
        T = g["Twc"].clone()
        s = be.GNSolver(prm, T, g["Xs"], g["Cs"], g["ii"], g["jj"], loc[0], loc[1], loc[2], e0, peers=peers)
        for it in range(6):
            if (rank + it + pattern) % world == 0:
                torch.cuda._sleep(600_000)
            s.accumulate()
            if (rank + it + pattern) % world == world - 1:
                torch.cuda._sleep(600_000)
            s.solve_step(0.0)
        torch.cuda.synchronize()
        if not torch.equal(T, T1):
            print(f"rank {rank}: N={N} {mode} skewed p2p pattern {pattern}: maxdiff={float((T - T1).abs().max()):.3e}",
                  flush=True)
            ok = False
    # packed edge records (FactorGraph's storage) through the same two exchanges: same bits again
    pk = be.edge_pack(loc[0], loc[1], loc[2])
    for p2p in (False, True):
        T = g["Twc"].clone()
        be.gauss_newton_sharded(prm, T, g["Xs"], g["Cs"], g["ii"], g["jj"], pk, None, None, e0, 6, 0.0,
                                edge_counts=counts, p2p=p2p)
        torch.cuda.synchronize()
        if not torch.equal(T, T1):
            print(f"rank {rank}: N={N} {mode} packed p2p={p2p}: maxdiff={float((T - T1).abs().max()):.3e}", flush=True)
            ok = False
    # early exit inside a sharded solve (delta_thresh huge => done after iteration 1, the remaining
    # epochs are no-ops that still signal and wait), immediately followed by another solve on the
    # same buffers with NO host synchronisation in between: the epoch/parity protocol must hold
    Ta, Tb = g["Twc"].clone(), g["Twc"].clone()
    be.gauss_newton_sharded(prm, Ta, g["Xs"], g["Cs"], g["ii"], g["jj"], loc[0], loc[1], loc[2], e0, 4, 1e9,
                            edge_counts=counts, p2p=True, check=False)
    be.gauss_newton_sharded(prm, Tb, g["Xs"], g["Cs"], g["ii"], g["jj"], loc[0], loc[1], loc[2], e0, 6, 0.0,
                            edge_counts=counts, p2p=True)
    torch.cuda.synchronize()
    T_one = g["Twc"].clone()
    be._gauss_newton(prm, T_one, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], 1, 0.0)
    if not (torch.equal(Ta, T_one) and torch.equal(Tb, T1)):
        print(f"rank {rank}: N={N} {mode} early exit then solve: {float((Ta - T_one).abs().max()):.3e} "
              f"{float((Tb - T1).abs().max()):.3e}", flush=True)
        ok = False

# BASELINE config 5 at reduced image size ("k-GPU dx equals 1-GPU dx"): 512 keyframes, 4096
# edges = 8192 directed, every edge's streams seeded by its GLOBAL id, so that each rank generating only
# its slice solves the very problem the single GPU solves
H5, W5, N5 = 48, 64, 512
K5 = synth.intrinsics(H5, W5, dev)
T_gt = synth.make_poses(N5, 1238, dev)
Xs5 = synth.render_pointmaps(T_gt, H5, W5, K5)
Cs5 = (1.0 + torch.exp(0.5 + 0.5 * synth.smooth_noise((N5, 1, H5, W5), 4.0 * W5 / synth.W0, synth._gen(1245, dev), dev)
                       .view(N5, H5 * W5, 1))).contiguous()
cnts = {k: None for k in range(1, 8)}
cnts["random"] = 4096 - sum(N5 - k for k in range(1, 8))
ii_u, jj_u = synth.make_edges(N5, cnts, 1238)
ii_d, jj_d = torch.cat([ii_u, jj_u]), torch.cat([jj_u, ii_u])
M5 = ii_d.shape[0]
ids5 = torch.arange(M5)
e0, cnt, counts = be.shard_edges(M5, world, rank)
idx_l, val_l, Q_l = synth.edge_streams(T_gt, Xs5, ii_d[e0:e0 + cnt], jj_d[e0:e0 + cnt], H5, W5, K5, 1251,
                                       edge_ids=ids5[e0:e0 + cnt])
idx_a, val_a, Q_a = synth.edge_streams(T_gt, Xs5, ii_d, jj_d, H5, W5, K5, 1251, edge_ids=ids5)
T0 = synth.perturb_poses(T_gt, 1267)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
ii_g, jj_g = ii_d.to(dev), jj_d.to(dev)
T_one = T0.clone()
(dx_one,) = be._gauss_newton(prm, T_one, Xs5, Cs5, ii_g, jj_g, idx_a, val_a, Q_a, 3, 0.0)
for p2p in (True, False):
    T = T0.clone()
    (dx_sh,) = be.gauss_newton_sharded(prm, T, Xs5, Cs5, ii_g, jj_g, idx_l, val_l, Q_l, e0, 3, 0.0, edge_counts=counts,
                                       p2p=p2p)
    torch.cuda.synchronize()
    if not (torch.equal(T, T_one) and torch.equal(dx_sh, dx_one)) or torch.equal(T, T0):
        print(f"rank {rank}: config 5 (N=512, {M5} edges, {H5}x{W5}) p2p={p2p}: pose maxdiff "
              f"{float((T - T_one).abs().max()):.3e}", flush=True)
        ok = False
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if int(flag.item()) == 1:
    print("MGPU_OK", flush=True)
dist.destroy_process_group()
