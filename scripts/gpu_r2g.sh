#!/bin/bash
mkdir -p gpurun_out
RSP_DEBUG_ERRORS=1 timeout 600 python -m pytest tests -m gpu -q -s -x -k "generator_many and 3-9" > gpurun_out/pytest_dbg_g.log 2>&1; echo "rc=$?"
grep -n "rsp\]" gpurun_out/pytest_dbg_g.log | head
