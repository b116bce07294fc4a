#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_chain.py tests/test_wire.py tests/test_gpu_doppler.py -m gpu -x -q -k "binary_points or binpoints or blocks_alone or jtag or fast_kernel or host" 2>&1 | tail -12 | tee gpurun_out/pytest_bp.txt
for k in app12 app12_14; do timeout 300 python scripts/bench_one.py $k 2>&1 | tail -1 | cut -c1-250; done | tee gpurun_out/app12b.txt
