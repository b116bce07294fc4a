#!/bin/bash
# weak scaling at 4 and 8 GPUs of one box (run under `gpurun --gpus 8`); the 1- and 2-GPU records come from the single-GPU visits
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi_8gpu.txt 2>&1
for n in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --no-cpu-baseline --no-configs \
    > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err
  echo "N=$n rc=$?"; tail -n 1 gpurun_out/bench_${n}gpu.json | cut -c1-160
done
