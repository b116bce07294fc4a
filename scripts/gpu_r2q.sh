#!/bin/bash
# full GPU suite + ncu captures of the fp32 kernel (C3 x 16 and the C5 shape) after the register-budget / 32-points-per-thread changes
TAG=${1:-q}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu_${TAG}.txt
cap() { # name kernel-regex command...
  local name=$1 rx=$2; shift 2
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 3 -c 1 -f -o gpurun_out/${name}_${TAG} "$@" > gpurun_out/ncu_${name}_${TAG}.log 2>&1; echo "ncu $name rc=$?"
  ncu -i gpurun_out/${name}_${TAG}.ncu-rep --page raw --csv > gpurun_out/${name}_${TAG}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${name}_${TAG}.ncu-rep --page details > gpurun_out/${name}_${TAG}_details.txt 2>/dev/null
  ncu -i gpurun_out/${name}_${TAG}.ncu-rep --page source --csv > gpurun_out/${name}_${TAG}_source.csv 2>/dev/null
  rm -f gpurun_out/${name}_${TAG}.ncu-rep
}
cap f32c3 chain_f32_fast_kernel python scripts/bench_one.py c3
cap f32c5 chain_f32_fast_kernel python scripts/bench_one.py c5f
