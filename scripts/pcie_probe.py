#!/usr/bin/env python
"""Raw pinned-host <-> device copy bandwidth of every rank at once (torchrun), both directions busy: the ceiling of bench.py's
`e2e` at N GPUs.  Prints one line per rank and the sum."""
import os, sys, time
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 256 << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory(); h_in.fill_(1)
h_out = torch.empty(n, dtype=torch.uint8).pin_memory(); h_out.fill_(2)
d_a, d_b = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.zeros(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def step():
    with torch.cuda.stream(s1): d_a.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2): h_out.copy_(d_b, non_blocking=True)
for _ in range(3): step()
torch.cuda.synchronize()
if world > 1: dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
K = 10
for _ in range(K): step()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
gbs = K * n / dt / 1e9
numa = "?"
try:
    bus = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
except Exception:
    bus = None
t = torch.tensor([gbs], device="cuda")
if world > 1:
    dist.all_reduce(t)
print(f"rank {rank}: {gbs:.1f} GB/s each way (both directions busy), cpus {sorted(os.sched_getaffinity(0))[:4]}..x{len(os.sched_getaffinity(0))}", flush=True)
if rank == 0:
    time.sleep(0.5)
    print(f"sum over {world} ranks: {t.item():.1f} GB/s each way", flush=True)
if world > 1: dist.destroy_process_group()
