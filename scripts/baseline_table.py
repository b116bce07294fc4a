#!/usr/bin/env python
"""Rewrites the results table of BASELINE.md section 5 from profiles/r02_<tag>_bench.json and ..._bench_ref.json:
  python scripts/baseline_table.py <tag>"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
PEAK = 6535.1


def last_json(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


line = last_json(os.path.join(ROOT, "profiles", f"r02_{tag}_bench.json"))
ref = last_json(os.path.join(ROOT, "profiles", f"r02_{tag}_bench_ref.json"))
f = lambda x: f"{x:,.0f}"  # noqa: E731
rows = []
v = line["value"]
rows.append(f"| **C5 headline (`bench.py value`): 1 frame = 8 RX x 4096 chirps x 16384-pt fixed CA (W64 G8) per GPU per step** | `chain_fixed_fast<14>` | 1 | {f(v)} | "
            f"{v * 8 / 1e3:.0f} | {v * 8 / 1e3 / 80:.1f} | {v * 8 / 1e3 / PEAK * 100:.1f} | {line['cpu_baseline']['value']:.1f} ({line['cpu_baseline']['cores']} thr) |")
e = line["e2e"]
rows.append(f"| same through `rsp_run_host` (`e2e`, 2 GiB H2D + 2 GiB D2H per step) | 3-deep chunk pipeline | 1 | {f(e['value'])} | {e['value'] * 4 / 1e3:.1f} each way over PCIe | - | - | - |")
d = e["detections_only"]
rows.append(f"| same, detections only (`e2e.detections_only`: `h_out = NULL`, {d['detections_per_step']:,} detections back) | 3-deep chunk pipeline | 1 | {f(d['value'])} | "
            f"{d['value'] * 4 / 1e3:.1f} H2D | - | - | - |")
two = os.path.join(ROOT, "profiles", f"r02_{tag}_bench_2gpu.json")
if os.path.exists(two):
    t2 = last_json(two)
    rows.append(f"| C5, 2 frames on 2 GPUs (`gpurun --gpus 2`, one frame per GPU per step) | `chain_fixed_fast<14>` | 2 | {f(t2['value'])} (e2e {f(t2['e2e']['value'])}) | "
                f"{t2['value'] * 8 / 1e3:.0f} | - | {t2['value'] * 8 / 1e3 / PEAK / 2 * 100:.1f} per GPU | - |")
for c in line["configs"]:
    one = f" (one frame {f(c['one_frame_Msamples_per_s'])})" if c.get("one_frame_Msamples_per_s") else ""
    g = c["GBps_algorithmic"]
    rows.append(f"| {c['config']} | `{c['variant']}` | 1 | {f(c['Msamples_per_s'])}{one} | {g:.0f} ({c['bytes_per_sample']} B/sample) | {g / 80:.1f} | {g / PEAK * 100:.1f} | - |")

path = os.path.join(ROOT, "BASELINE.md")
s = open(path).read()
i0 = s.index("| **C5 headline")
i1 = s.index("\nReference arm (`bench.py --impl reference`")
s = s[:i0] + "\n".join(rows) + "\n" + s[i1:]
s = re.sub(r"[0-9.]+ Msamples/s on 16384-point rows of the same workload; `e2e` ÷ that = \d+ ×, device-resident\n`value` ÷ that = \d+ ×",
           f"{ref['value']:.1f} Msamples/s on 16384-point rows of the same workload; `e2e` ÷ that = {e['value'] / ref['value']:.0f} ×, device-resident\n"
           f"`value` ÷ that = {v / ref['value']:.0f} ×", s)
s = re.sub(r"\(`rsp_gather_detections`, outside the step\): [0-9.]+ ms at N = 1", f"(`rsp_gather_detections`, outside the step): {line['gather_ms']:.3f} ms at N = 1", s)
if os.path.exists(two):
    s = re.sub(r"ms at N = 1, [0-9.]+ ms at N = 2", f"ms at N = 1, {last_json(two)['gather_ms']:.3f} ms at N = 2", s)
open(path, "w").write(s)
print(len(rows), "rows;", "value", f(v), "e2e", f(e["value"]), "ref", f"{ref['value']:.1f}")
