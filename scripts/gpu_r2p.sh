#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_round2.py -m gpu -x -q -k "gos or os" 2>&1 | tail -8 | tee gpurun_out/pytest_sparse.txt
for k in c4os16 c4osb c4os16d c4osbd; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-170; done | tee gpurun_out/sparse_ab.txt
