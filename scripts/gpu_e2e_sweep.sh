#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2 3; do for cfg in "3 16" "3 32" "3 64"; do set -- $cfg; RSP_PIPE_DEPTH=$1 RSP_PIPE_CHUNK_MIB=$2 timeout 120 python scripts/e2e_sweep.py 2>&1 | tail -1; done; done | tee gpurun_out/e2e_sweep_o.txt
