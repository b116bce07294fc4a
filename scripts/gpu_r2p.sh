#!/bin/bash
mkdir -p gpurun_out
P=$PWD/rsp_chains_b200
timeout 900 python -m pytest tests -m gpu -x -q -k "f32 or fp32" 2>&1 | tail -5 | tee gpurun_out/pytest_f32_t.txt
for k in c3 fp10 fp12 fp13 fp14 c5f; do for v in "" _f32nopf; do echo "lib$v $k:"; RSP_B200_LIB=$P/librsp_b200$v.so timeout 200 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-200; done; done | tee gpurun_out/f32_variants_t.txt
