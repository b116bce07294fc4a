#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/latency_probe.py 2>&1 | tail -1 | tee gpurun_out/latency_probe2.txt
