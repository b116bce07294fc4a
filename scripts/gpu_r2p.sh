#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py -m gpu -x -q -k "peak_grouping or fast_kernel or generator" 2>&1 | tail -8 | tee gpurun_out/pytest_pg.txt
