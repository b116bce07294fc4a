#!/bin/bash
# round-2 GPU visit: parity tests, smoke, bench (+ per-config lines), reference arm, ncu launch list + full capture of the headline kernel with
# the per-instruction source page.  $1 = tag.  Outputs -> gpurun_out/
TAG=${1:-a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi_${TAG}.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_${TAG}.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${TAG}.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke_${TAG}.log
timeout 900 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench rc=$?"
cat gpurun_out/bench_${TAG}.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG}.csv python bench.py --steps 2 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_list_${TAG}.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chain_fixed_fast -s 3 -c 1 -f -o gpurun_out/fx14_full_${TAG} python bench.py --steps 2 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_full_${TAG}.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/fx14_full_${TAG}.ncu-rep --page raw --csv > gpurun_out/fx14_full_${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/fx14_full_${TAG}.ncu-rep --page source --csv > gpurun_out/fx14_full_${TAG}_source.csv 2>/dev/null
tail -n 6 gpurun_out/pytest_gpu_${TAG}.log gpurun_out/smoke_${TAG}.log
# A/B: N = 16384 with 16 points per thread (one 1024-thread CTA per SM, cp.async staging) against the default 32 points per thread
for lp in 4 5; do RSP_FAST_LP14=$lp timeout 300 python bench.py --steps 10 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-configs 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('LP14=$lp', round(d['value']), 'Msamples/s frac', round(d['roofline']['frac'],4))" | tee -a gpurun_out/lp14_ab_${TAG}.txt; done
