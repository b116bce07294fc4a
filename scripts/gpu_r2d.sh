#!/bin/bash
TAG=${1:-d}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 25 gpurun_out/pytest_gpu_${TAG}.log
for l2 in 0 1; do for k in fx14 c2 fx12; do echo "l2_prefetch=$l2 $k:"; RSP_FAST_L2_PREFETCH=$l2 timeout 200 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-200; done; done | tee -a gpurun_out/l2prefetch_${TAG}.txt
