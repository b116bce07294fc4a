#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "wide or windowing or overflow" 2>&1 | tail -8 | tee gpurun_out/pytest_wide.txt
timeout 300 python scripts/wide_probe.py 2>&1 | tail -4 | tee gpurun_out/wide_probe2.txt
echo "RSP_WIDE_PASSES=0"; RSP_WIDE_PASSES=0 timeout 300 python scripts/wide_probe.py 2>&1 | tail -4 | tee gpurun_out/wide_probe0.txt
