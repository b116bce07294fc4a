#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py -m gpu -x -q -k "bit_plane or gos or os or binary_points" 2>&1 | tail -12 | tee gpurun_out/pytest_bpg.txt
for k in c4os16 c4osb; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-170; done | tee gpurun_out/bpg_bench.txt
