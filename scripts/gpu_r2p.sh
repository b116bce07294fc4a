#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "gos or os_cfar" 2>&1 | tail -8 | tee gpurun_out/pytest_gosb.txt
for v in 1 0; do echo "RSP_FAST_GOS_BITPLANES=$v"; RSP_FAST_GOS_BITPLANES=$v timeout 300 python scripts/bench_one.py c4os16 2>&1 | tail -1 | cut -c1-200; done | tee gpurun_out/gosb_ab.txt
