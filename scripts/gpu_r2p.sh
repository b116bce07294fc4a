#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py tests/test_gpu_f32_gen.py -m gpu -x -q -k "binary_points or binpoints or fast_kernel or generator or host or c2 or c5" 2>&1 | tail -12 | tee gpurun_out/pytest_bpc.txt
for k in app12 c5 c2; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-200; done | tee gpurun_out/bpc_bench.txt
