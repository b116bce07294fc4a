#!/bin/bash
# ncu evidence for the kernels added late in round 2: reading (B) bit planes, the binary-point CA kernel (app parameters), 256-point CA kernel
mkdir -p gpurun_out
cap() { # name kernel-regex command...
  local name=$1 rx=$2; shift 2
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 3 -c 1 -f -o gpurun_out/${name}_zz "$@" > gpurun_out/ncu_${name}_zz.log 2>&1; echo "ncu $name rc=$?"
  ncu -i gpurun_out/${name}_zz.ncu-rep --page raw --csv > gpurun_out/${name}_zz_raw.csv 2>/dev/null
  ncu -i gpurun_out/${name}_zz.ncu-rep --page source --csv > gpurun_out/${name}_zz_source.csv 2>/dev/null
  rm -f gpurun_out/${name}_zz.ncu-rep
}
cap osb12 chain_fixed_fast_os_kernel python scripts/bench_one.py c4osb
cap app12 chain_fixed_fast_kernel python scripts/bench_one.py app12
cap fx8 chain_fixed_fast_kernel python scripts/bench_one.py small8
