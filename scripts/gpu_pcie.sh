#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
{ nvidia-smi topo -m; lscpu | grep -i "numa\|socket\|model name\|^CPU(s)"; for d in /sys/bus/pci/devices/*; do c=$(cat $d/class 2>/dev/null); [ "$c" = "0x030200" ] && echo "$d numa=$(cat $d/numa_node) speed=$(cat $d/current_link_speed) width=$(cat $d/current_link_width)"; done; } > gpurun_out/topo_${N}gpu.txt 2>&1
python scripts/pcie_probe.py > gpurun_out/pcie_1.txt 2>&1
for n in 2 4 8; do [ $n -gt $N ] && break; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29519 scripts/pcie_probe.py > gpurun_out/pcie_$n.txt 2>&1; done
cat gpurun_out/topo_${N}gpu.txt; grep -h "rank\|sum" gpurun_out/pcie_*.txt
