#!/bin/bash
# round-2 visit c: parity tests + A/B of compile-time variants (scripts/build_variant.sh).  Outputs -> gpurun_out/
TAG=${1:-c}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 15 gpurun_out/pytest_gpu_${TAG}.log
P=$PWD/rsp_chains_b200
for v in "" _mb1 _twvol; do echo "lib$v fx14:"; RSP_B200_LIB=$P/librsp_b200$v.so timeout 200 python scripts/bench_one.py fx14 2>&1 | tail -1; done | tee -a gpurun_out/variants_${TAG}.txt
for v in "" _mb4s10; do echo "lib$v c2:"; RSP_B200_LIB=$P/librsp_b200$v.so timeout 200 python scripts/bench_one.py c2 2>&1 | tail -1; done | tee -a gpurun_out/variants_${TAG}.txt
