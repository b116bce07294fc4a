#!/bin/bash
TAG=${1:-f}
mkdir -p gpurun_out
RSP_DEBUG_ERRORS=1 timeout 1500 python -m pytest tests -m gpu -q -s -k "generator or reading_b" > gpurun_out/pytest_dbg_${TAG}.log 2>&1; echo "dbg pytest rc=$?"
grep -n "pending CUDA\|PASSED\|FAILED\|passed\|failed" gpurun_out/pytest_dbg_${TAG}.log | head -40
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 14 gpurun_out/pytest_gpu_${TAG}.log
