#!/bin/bash
TAG=${1:-k}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 6 gpurun_out/pytest_gpu_${TAG}.log
P=$PWD/rsp_chains_b200
for v in "" _old14; do echo "lib$v fx14:"; RSP_B200_LIB=$P/librsp_b200$v.so timeout 200 python scripts/bench_one.py fx14 2>&1 | tail -1 | cut -c1-190; done | tee -a gpurun_out/true_last_${TAG}.txt
for v in "" _old10; do echo "lib$v c2:"; RSP_B200_LIB=$P/librsp_b200$v.so timeout 200 python scripts/bench_one.py c2 2>&1 | tail -1 | cut -c1-190; done | tee -a gpurun_out/true_last_${TAG}.txt
for k in fx11 fx12 fx13; do echo "$k:"; timeout 200 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-190; done | tee -a gpurun_out/true_last_${TAG}.txt
