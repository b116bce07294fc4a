#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_doppler.py -m gpu -x -q -k "fp32_flavour" 2>&1 | tail -25 | tee gpurun_out/pytest_rdf32.txt
