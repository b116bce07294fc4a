#!/bin/bash
mkdir -p gpurun_out
{ bash scripts/gpu_ab.sh fx14; bash scripts/gpu_ab.sh c5; } 2>&1 | tee gpurun_out/tw_ab.txt
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "fast or c5 or window or all_sizes" > gpurun_out/pytest_fast.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_fast.log
