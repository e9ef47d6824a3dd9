"""Where an edge-sharded GN iteration goes (config 3 shape: N=32, 200 directed edges, 512x384): CUDA events after
accumulate and after solve_step of every iteration on every rank (peer-memory exchange), and the whole call.
Usage: torchrun --nproc-per-node N profiles/time_sharded_breakdown.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mast3r-slam_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from tests import synth  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
g = synth.make_graph(N=32, seed=1236, device=dev)
M = g["M"]
e0, cnt, counts = be.shard_edges(M, world, rank)
prm = be._params(be.GN_RAYS, 0.003, 10.0, 0.0, 1.5)
pk = be.edge_pack(g["idx"][e0:e0 + cnt].contiguous(), g["valid"][e0:e0 + cnt].contiguous(), g["Q"][e0:e0 + cnt].contiguous())
peers = be.PeerBlocks.get(M, dev) if world > 1 else None
iters = 10


def one(record):
    T = g["Twc"].clone()
    s = be.GNSolver(prm, T, g["Xs"], g["Cs"], g["ii"], g["jj"], pk, None, None, e0, peers=peers)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * iters + 1)]
    if record:
        ev[0].record()
    for it in range(iters):
        s.accumulate()
        if record:
            ev[2 * it + 1].record()
        s.solve_step(0.0)
        if record:
            ev[2 * it + 2].record()
    return ev


for _ in range(3):
    one(False)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
ev = one(True)
torch.cuda.synchronize()
acc = [ev[2 * i].elapsed_time(ev[2 * i + 1]) * 1e3 for i in range(iters)]
sol = [ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) * 1e3 for i in range(iters)]
a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
if world > 1:
    dist.barrier()
a.record()
T = g["Twc"].clone()
be.gauss_newton_sharded(prm, T, g["Xs"], g["Cs"], g["ii"], g["jj"], pk, None, None, e0, iters, 0.0, edge_counts=counts,
                        p2p=world > 1)
z.record()
torch.cuda.synchronize()
for r in range(world):
    if r == rank:
        print(f"rank {rank}/{world}: {cnt} edges; per iteration us: accumulate (launch to end) median {sorted(acc)[iters // 2]:.1f}, "
              f"solve_step (incl. waiting for the peers' flags) median {sorted(sol)[iters // 2]:.1f}; first three: "
              f"{[round(x) for x in acc[:3]]} / {[round(x) for x in sol[:3]]}; whole call {a.elapsed_time(z) * 1e3:.0f} us "
              f"= {a.elapsed_time(z) * 100:.1f} us per iteration", flush=True)
    if world > 1:
        dist.barrier()
if world > 1:
    dist.destroy_process_group()
