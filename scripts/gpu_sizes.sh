#!/bin/bash
# GPU visit: parity tests, then the fixed / fp32 chain at every size (512 MiB footprint) and the config sweep
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest_gpu.log
: > gpurun_out/sizes.jsonl
for k in ${SIZES:-fx10 fx11 fx12 fx13 fx14}; do timeout 300 python scripts/bench_one.py $k >> gpurun_out/sizes.jsonl 2>> gpurun_out/sizes.err; done
python - <<'PY'
import json
for l in open('gpurun_out/sizes.jsonl'):
    d=json.loads(l); print(d['config'][:40].ljust(40), d['variant'].ljust(22), d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])
PY
timeout 900 python scripts/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    d=json.loads(l); print(d['config'][:52].ljust(52), d['variant'].ljust(22), d['Msamples_per_s'], d['frac_of_measured_hbm_peak'])
PY
