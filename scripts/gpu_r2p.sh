#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py -m gpu -x -q -k "stored_profiles or gos or os or blocks_alone or reading_b" 2>&1 | tail -8 | tee gpurun_out/pytest_co.txt
for k in c4cfar c4cfarb; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-200; done | tee gpurun_out/cfar_only.txt
