#!/usr/bin/env python
"""BASELINE.md section 5 rows from one visit's artefacts: python scripts/results_table.py profiles/r01_l"""
import json, sys
pre = sys.argv[1]
peak = json.load(open(pre + "_bench.json"))["roofline"]["peak"]
cpu = {}
try:
    for l in open(pre + "_cpu_configs.jsonl"):
        d = json.loads(l); cpu[d["config"].split()[0]] = d
except FileNotFoundError:
    pass
print("| Config | Variant | GPUs | Msamples/s | GB/s achieved | %% of 8 TB/s | %% of measured %.1f GB/s | CPU 1-thread Msamples/s | CPU all-thread Msamples/s |" % peak)
print("|---|---|---|---|---|---|---|---|---|")
for l in open(pre + "_configs.jsonl"):
    d = json.loads(l)
    c = cpu.get(d["config"].split()[0])
    fixed = "fixed" in d["variant"]
    c1 = f'{c["Msamples_per_s_1_thread"]}' if c and fixed else "-"
    ca = f'{c["Msamples_per_s_all_threads"]} ({c["threads"]} thr)' if c and fixed else "-"
    print(f'| {d["config"]} | `{d["variant"]}` | 1 | {d["Msamples_per_s"]:,.0f} | {d["GBps_algorithmic"]:.0f} | {100 * d["GBps_algorithmic"] / 8000:.1f} | '
          f'{100 * d["frac_of_measured_hbm_peak"]:.1f} | {c1} | {ca} |')
