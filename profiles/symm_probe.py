"""Probe: torch symmetric memory (peer-mapped buffers over NVLink) on this box.
torchrun --nproc-per-node 2 profiles/symm_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm.empty((1024,), dtype=torch.float64, device=dev)
t.fill_(rank + 1)
h = symm.rendezvous(t, dist.group.WORLD)
print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in h.signal_pad_ptrs],
      "signal_pad_size", h.signal_pad_size, "attrs", [a for a in dir(h) if not a.startswith("_")], flush=True)
h.barrier()
peer = h.get_buffer((rank + 1) % world, (1024,), torch.float64)
print(rank, "peer value", float(peer[0]), flush=True)
peer[1] = 100 + rank          # remote store
h.barrier()
torch.cuda.synchronize()
print(rank, "local t[1] after peer store", float(t[1]), flush=True)
dist.destroy_process_group()
