#!/bin/bash
# two GPUs: the multi-GPU test (sharded rows + NCCL gather through the C ABI) and the bench at N = 2
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/pytest_multi_i.log 2>&1; echo "multi pytest rc=$?"; tail -5 gpurun_out/pytest_multi_i.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu_i.json 2> gpurun_out/bench_2gpu_i.err; echo "bench2 rc=$?"
head -c 300 gpurun_out/bench_2gpu_i.json; echo; tail -3 gpurun_out/bench_2gpu_i.err
python - <<'PY'
import json
s=open('gpurun_out/bench_2gpu_i.json').read()
d=json.loads(s[s.index('{'):])
print(round(d['value']), d['n_gpus'], round(d['e2e']['value']), d['gather_ms'], d['gather'][:120])
PY
