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
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if int(flag.item()) == 1:
    print("MGPU_OK", flush=True)
dist.destroy_process_group()
