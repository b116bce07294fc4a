#!/bin/bash
# round-2 visit b: parity tests + A/B of the start skew of the two CTAs per SM at N = 16384.  Outputs -> gpurun_out/
TAG=${1:-b}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 15 gpurun_out/pytest_gpu_${TAG}.log
for cfg in "5 0" "5 5000" "5 15000" "5 30000" "4 0"; do set -- $cfg
  RSP_FAST_LP14=$1 RSP_FAST_SKEW_NS=$2 timeout 300 python bench.py --steps 10 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('LP14=$1 skew=$2', round(d['value']), 'Msamples/s frac', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']), 'dets-only', round(d['e2e']['detections_only']['value']), d['e2e']['detections_only']['detections_per_step'])" | tee -a gpurun_out/skew_ab_${TAG}.txt; done
