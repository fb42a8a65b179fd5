"""2-GPU probe of torch symmetric memory on the box: rendezvous, peer pointers, P2P copy, barrier.
   torchrun --nproc-per-node 2 tools/probe_symm.py"""
import os
import time

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
n = 1 << 20
buf = symm.empty((world, n), dtype=torch.float64, device=dev)
buf.zero_()
hdl = symm.rendezvous(buf, dist.group.WORLD)
print(rank, "backend", getattr(hdl, "get_backend", lambda: "?")() if False else "", "ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast", hdl.has_multicast_support if not callable(hdl.has_multicast_support) else "fn", flush=True)
peers = [hdl.get_buffer(r, buf.shape, buf.dtype) for r in range(world)]
src = torch.full((n,), float(rank + 1), dtype=torch.float64, device=dev)
hdl.barrier(channel=0)
t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
t0.record()
for it in range(20):
    for r in range(world):
        peers[r][rank].copy_(src, non_blocking=True)
    hdl.barrier(channel=0)
t1.record()
torch.cuda.synchronize()
ok = all(bool((buf[r] == r + 1).all()) for r in range(world))
print(rank, "gathered ok", ok, "ms per exchange of", n * 8 / 1e6, "MB slots:", t0.elapsed_time(t1) / 20, flush=True)
dist.barrier()
dist.destroy_process_group()
