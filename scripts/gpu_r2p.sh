#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -12 | tee gpurun_out/smoke_v.txt
timeout 600 python - <<'PY' 2>&1 | tail -8 | tee gpurun_out/doppler_lines_v.txt
import json, torch, bench
torch.cuda.set_device(0)
for l in bench.doppler_lines(torch, 6535.1): print(json.dumps(l))
PY
