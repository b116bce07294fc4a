#!/bin/bash
TAG=${1:-e}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 12 gpurun_out/pytest_gpu_${TAG}.log
for lp in 5 4; do echo "LP14=$lp fx14:"; RSP_FAST_LP14=$lp timeout 200 python scripts/bench_one.py fx14 2>&1 | tail -1 | cut -c1-200; done | tee -a gpurun_out/lp_l2_${TAG}.txt
for f in 1 0; do echo "RSP_GEN_FUSED=$f:"; RSP_GEN_FUSED=$f timeout 300 python -c "
import json, torch, bench
torch.cuda.set_device(0)
for l in bench.generator_lines(torch, bench.measured_peak()[0]): print(json.dumps(l))
"; done 2>&1 | tee -a gpurun_out/gen_fused_${TAG}.txt
