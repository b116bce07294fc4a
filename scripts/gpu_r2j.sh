#!/bin/bash
# round-2 visit j: final-state artefacts: bench, reference arm, launch list, full ncu captures (raw + per-instruction source page) of the headline
# kernel (N = 16384), the N = 1024 kernel, the fp32 kernel and the GOS kernel
TAG=${1:-j}
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG}.csv python bench.py --steps 2 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_list_${TAG}.log 2>&1; echo "ncu list rc=$?"
cap() { # name kernel-regex command...
  local name=$1 rx=$2; shift 2
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 3 -c 1 -f -o gpurun_out/${name}_${TAG} "$@" > gpurun_out/ncu_${name}_${TAG}.log 2>&1; echo "ncu $name rc=$?"
  ncu -i gpurun_out/${name}_${TAG}.ncu-rep --page raw --csv > gpurun_out/${name}_${TAG}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${name}_${TAG}.ncu-rep --page source --csv > gpurun_out/${name}_${TAG}_source.csv 2>/dev/null
  rm -f gpurun_out/${name}_${TAG}.ncu-rep
}
cap fx14 chain_fixed_fast_kernel python bench.py --steps 2 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs
cap fx10 chain_fixed_fast_kernel python scripts/bench_one.py c2
cap f32c3 chain_f32_fast_kernel python scripts/bench_one.py c3
cap gos12 chain_fixed_fast_gosb_kernel python scripts/bench_one.py c4os16
ls -la gpurun_out | tail -20
