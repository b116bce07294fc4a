#!/bin/bash
# GPU visit for the OS-CFAR kernel: parity tests, C4 timings
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "gos or lis or GOS" > gpurun_out/pytest_gos.log 2>&1; echo "pytest gos rc=$?"; tail -n 8 gpurun_out/pytest_gos.log
timeout 300 python scripts/bench_one.py c4os > gpurun_out/c4os.json 2> gpurun_out/c4os.err; echo "c4os rc=$?"; cat gpurun_out/c4os.json; tail -n 3 gpurun_out/c4os.err
timeout 300 python scripts/bench_one.py c4os16 > gpurun_out/c4os16.json 2> gpurun_out/c4os16.err; echo "c4os16 rc=$?"; cat gpurun_out/c4os16.json; tail -n 3 gpurun_out/c4os16.err
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_gpu.log
