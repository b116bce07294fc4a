#!/bin/bash
# full GPU suite + smoke, then the final-state artefacts (scripts/gpu_r2j.sh <tag>)
TAG=${1:-y}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu_$TAG.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -10 | tee gpurun_out/smoke_$TAG.txt
bash scripts/gpu_r2j.sh $TAG
