#!/bin/bash
# quick GPU visit: parity tests + all-config sweep (no bench.py)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/pytest_gpu.log
timeout 900 python scripts/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    d=json.loads(l); print(d['config'][:52].ljust(52), d['variant'].ljust(22), d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])
PY
tail -n 3 gpurun_out/configs.err
