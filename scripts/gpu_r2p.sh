#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_f32_gen.py tests/test_zz_no_alloc.py -m gpu -x -q -k "generator or nco or rasterized or static" 2>&1 | tail -8 | tee gpurun_out/pytest_gt.txt
for v in 1 0; do RSP_GEN_TABLE=$v timeout 300 python - <<'PY' 2>&1 | tail -2 | cut -c1-220
import json, torch, bench
torch.cuda.set_device(0)
for l in bench.generator_lines(torch, 6535.1): print(json.dumps({k: l[k] for k in ("config", "variant", "ms", "Msamples_per_s", "frac")}))
PY
done | tee gpurun_out/gen_table_ab.txt
