#!/bin/bash
mkdir -p gpurun_out
echo "== alone" | tee gpurun_out/pytest_noalloc.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -12 | tee -a gpurun_out/pytest_noalloc.txt
echo "== full suite" | tee -a gpurun_out/pytest_noalloc.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee -a gpurun_out/pytest_noalloc.txt
