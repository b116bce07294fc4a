#!/bin/bash
# A/B of the cp.async staging of the fast fixed kernel (RSP_FAST_STAGE=0/1) over the transform sizes, then the fast-path parity tests
mkdir -p gpurun_out
for k in c2 fx10 fx11 fx12 fx13 fx14; do
  for s in 0 1; do
    RSP_FAST_STAGE=$s python scripts/bench_one.py $k 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$k stage=$s', d['variant'], d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])"
  done
done | tee gpurun_out/stage_ab.txt
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "fast or c2 or c5 or window or all_sizes or run_host" > gpurun_out/pytest_fast.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_fast.log
