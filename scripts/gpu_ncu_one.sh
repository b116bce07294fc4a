#!/bin/bash
# ncu --set full of one launch of the chain kernel of one configuration: $1 = key of scripts/bench_one.py, $2 = tag
KEY=${1:-c3}; TAG=${2:-x}
mkdir -p gpurun_out
timeout 300 python scripts/bench_one.py $KEY > gpurun_out/one_${TAG}.json 2> gpurun_out/one_${TAG}.err; echo "run rc=$?"; cat gpurun_out/one_${TAG}.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chain_ -s 4 -c 1 -f -o gpurun_out/full_${TAG} python scripts/bench_one.py $KEY > gpurun_out/ncu_${TAG}.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/full_${TAG}.ncu-rep --page raw --csv > gpurun_out/full_${TAG}_raw.csv 2>/dev/null
