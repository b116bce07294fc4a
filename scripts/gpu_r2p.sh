#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "below_1024" 2>&1 | tail -12 | tee gpurun_out/pytest_small2.txt
