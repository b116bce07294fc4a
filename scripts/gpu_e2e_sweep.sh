#!/bin/bash
mkdir -p gpurun_out
for d in 2 3 4 6; do for c in 2 4 8 16 32 64; do RSP_PIPE_DEPTH=$d RSP_PIPE_CHUNK_MIB=$c timeout 120 python scripts/e2e_sweep.py 2>&1 | tail -1; done; done | tee gpurun_out/e2e_sweep.txt
nvidia-smi -q | grep -A3 "PCI" | grep -iE "gen|width" | head -8
python - <<'PY'
import torch,time
a=torch.empty(1<<28,dtype=torch.uint8).pin_memory(); b=torch.empty(1<<28,dtype=torch.uint8,device='cuda'); c=torch.empty(1<<28,dtype=torch.uint8).pin_memory(); d=torch.empty(1<<28,dtype=torch.uint8,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
for name,fn in (("h2d",lambda: b.copy_(a,non_blocking=True)),("d2h",lambda: c.copy_(d,non_blocking=True))):
    fn(); torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(8): fn()
    torch.cuda.synchronize(); print(name, 8*(1<<28)/(time.perf_counter()-t)/1e9,"GB/s")
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(8):
    with torch.cuda.stream(s1): b.copy_(a,non_blocking=True)
    with torch.cuda.stream(s2): c.copy_(d,non_blocking=True)
torch.cuda.synchronize(); print("both", 8*(1<<28)/(time.perf_counter()-t)/1e9,"GB/s each way")
PY
