#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py -m gpu -x -q -k "below_1024 or binary_points or peak_grouping or binpoints" 2>&1 | tail -12 | tee gpurun_out/pytest_pgbp.txt
for k in app12 app12_14; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-200; done | tee gpurun_out/app12c.txt
