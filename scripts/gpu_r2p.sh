#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "os or reading_b" 2>&1 | tail -8 | tee gpurun_out/pytest_os_z2.txt
timeout 300 python scripts/bench_one.py c4osb 2>&1 | tail -1 | cut -c1-200 | tee gpurun_out/osb_bench2.txt
