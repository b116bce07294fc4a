#!/usr/bin/env python
"""CPU baseline per BASELINE.json configuration: the oracle port (oracle/rsp_oracle.c, the integer restatement of the reference's
arithmetic; the Scala/Chisel reference itself cannot run without a JVM) on 1 thread and on all host threads, bounded samples.
`python scripts/cpu_configs.py` -> one JSON line per configuration.  Test/benchmark infrastructure, never used by the product."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
from oracle import oracle as O
from oracle.glue import oracle_cfgs
from rsp_chains_b200 import synth
from rsp_chains_b200.params import GOSCFARType
import bench_configs as B

O.build()
threads = max(1, len(os.sched_getaffinity(0)))
ca = lambda n, W, G, ds, mode=0: dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=ds, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)
os_regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
cases = [("C1 1 x 1024 fixed CA (W32 G4)", B.fixed_params(1024, 64, 4), 10, ca(1024, 32, 4, 5), 1),
         ("C2 256 x 1024 fixed CA (W16 G4)", B.fixed_params(1024, 64, 4), 10, ca(1024, 16, 4, 4), 256),
         ("C3 shape 2048-pt in fixed point, GO (W32 G4)", B.fixed_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, 1), 256),
         ("C4 4096-pt fixed GOS (64 cells per side, k=48)", B.fixed_params(4096, 64, 4, GOSCFARType), 12, os_regs, 64),
         ("C5 16384-pt fixed CA (W64 G8)", B.fixed_params(16384, 64, 8), 14, ca(16384, 64, 8, 6), 64)]
for name, params, log_n, regs, rows in cases:
    n = 1 << log_n
    fft, mag, cfar = oracle_cfgs(O, params, log_n, 2, regs)
    beats = synth.chirps_fixed(rows, n, seed=1)
    res = {}
    for nt in (1, threads):
        O.chain_fixed_words(fft, mag, cfar, beats[:max(1, rows // 8)], n_threads=nt)
        reps, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 2.0 or reps < 2:
            O.chain_fixed_words(fft, mag, cfar, beats, n_threads=nt); reps += 1
        res[nt] = rows * n * reps / (time.perf_counter() - t0) / 1e6
    print(json.dumps({"config": name, "rows_per_call": rows, "n": n, "Msamples_per_s_1_thread": round(res[1], 2),
                      "Msamples_per_s_all_threads": round(res[threads], 2), "threads": threads, "kind": "port (oracle/rsp_oracle.c)"}), flush=True)
