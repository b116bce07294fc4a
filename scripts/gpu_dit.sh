#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "radix2 or radix22" > gpurun_out/pytest_dit.log 2>&1; echo "dit rc=$?"; tail -n 12 gpurun_out/pytest_dit.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log
