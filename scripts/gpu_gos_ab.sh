#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "gos or os_cfar or cfar_cores" > gpurun_out/pytest_gos.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_gos.log
{ bash scripts/gpu_ab.sh c4os; bash scripts/gpu_ab.sh c4os16; } 2>&1 | tee gpurun_out/gos_ab.txt
