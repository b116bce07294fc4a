#!/usr/bin/env python
"""Per-call cost of rsp_run on one 1024-point chirp (BASELINE configs[0]): back-to-back calls through ctypes, wall clock and CUDA events,
next to an empty torch kernel launch for scale."""
import sys, os, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
import rsp_chains_b200 as R
from rsp_chains_b200 import synth
torch.cuda.set_device(0)
n = 1024
regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
d_in = torch.from_numpy(synth.chirps_fixed(1, n, seed=1).view(np.int32)).cuda()
d_out = torch.empty_like(d_in)
res = {}
with R.FftMagCfarChainVanilla(bench.fixed_params(n, 64, 4)) as ch:
    ch.configure(10, 2, regs)
    for reps in (2000,):
        for _ in range(200): ch.run(d_in, 1, 1, out=d_out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        for _ in range(reps): ch.run(d_in, 1, 1, out=d_out)
        e1.record(); t_issue = time.perf_counter() - t0
        torch.cuda.synchronize(); t_all = time.perf_counter() - t0
        res["rsp_run_us_per_call_device"] = e0.elapsed_time(e1) * 1e3 / reps
        res["rsp_run_us_per_call_host_issue"] = t_issue * 1e6 / reps
        res["rsp_run_us_per_call_wall"] = t_all * 1e6 / reps
with R.FftMagCfarChainVanilla(bench.fixed_params(n, 64, 4)) as ch:      # the same calls captured once in a CUDA graph of 100 launches and replayed
    ch.configure(10, 2, regs)
    ch.run(d_in, 1, 1, out=d_out); torch.cuda.synchronize()
    g, st = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.graph(g, stream=st):
        for _ in range(100): ch.run(d_in, 1, 1, out=d_out, stream=st)
    for _ in range(5): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): g.replay()
    e1.record(); torch.cuda.synchronize()
    res["graph_replay_us_per_call_device"] = e0.elapsed_time(e1) * 1e3 / 2000
x = torch.zeros(1, device="cuda")
for _ in range(200): x.add_(1)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(2000): x.add_(1)
t_issue = time.perf_counter() - t0; torch.cuda.synchronize()
res["torch_add_us_per_call_host_issue"] = t_issue * 1e6 / 2000
res["torch_add_us_per_call_wall"] = (time.perf_counter() - t0) * 1e6 / 2000
print(json.dumps(res))
