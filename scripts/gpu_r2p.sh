#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_doppler.py tests/test_gpu_chain.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_fftfast_x.txt
for v in 1 0; do echo "RSP_FAST_FFT=$v"; RSP_FAST_FFT=$v timeout 600 python - <<'PY' 2>&1 | tail -4 | cut -c1-260
import json, torch, bench
torch.cuda.set_device(0)
for l in bench.doppler_lines(torch, 6535.1): print(json.dumps({k: l[k] for k in ("config", "ms", "Msamples_per_s", "frac")}))
PY
done | tee gpurun_out/doppler_lines_x.txt
