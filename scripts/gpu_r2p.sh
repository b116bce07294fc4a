#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "os or reading_b" 2>&1 | tail -15 | tee gpurun_out/pytest_os_z.txt
