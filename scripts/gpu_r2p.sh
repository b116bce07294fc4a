#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_f32_gen.py -m gpu -x -q -k "f32 or fp32" 2>&1 | tail -12 | tee gpurun_out/pytest_pgf.txt
timeout 300 python scripts/bench_one.py c3 2>&1 | tail -1 | cut -c1-170 | tee gpurun_out/pgf_bench.txt
