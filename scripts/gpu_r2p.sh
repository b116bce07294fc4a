#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_f32_gen.py -m gpu -x -q -k "rasterized or generator or nco" 2>&1 | tail -8 | tee gpurun_out/pytest_raster.txt
