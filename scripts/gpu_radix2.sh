#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "radix2 or radix22" > gpurun_out/pytest_radix2.log 2>&1; echo "radix2 rc=$?"
tail -n 30 gpurun_out/pytest_radix2.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/pytest_gpu.log
