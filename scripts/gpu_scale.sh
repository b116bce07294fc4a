#!/bin/bash
# scaling bench on the GPUs of one box: bash scripts/gpu_scale.sh N  (run under `gpurun --gpus N`).  Outputs -> gpurun_out/
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi_${N}gpu.txt 2>&1
for n in 1 2 4 8; do
  [ $n -gt $N ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --no-cpu-baseline \
      > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err
  fi
  echo "N=$n rc=$?"; tail -n 1 gpurun_out/bench_${n}gpu.json | cut -c1-200
done
