#!/bin/bash
TAG=${1:-h}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
tail -n 14 gpurun_out/pytest_gpu_${TAG}.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${TAG}.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke_${TAG}.log
timeout 900 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_${TAG}.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_'"$TAG"'.json"))
print(round(d['value']), d['roofline']['frac'], d['e2e']['value'], d['e2e']['detections_only'], d['gather_ms'], d['gather'][:80])
for c in d.get('configs', []): print(c.get('config','?')[:70], c.get('Msamples_per_s'), c.get('frac'), c.get('variant'))
PY
