#!/usr/bin/env python
"""Throughput of the wide FFT block (24-bit data and twiddles, Convergent; with and without grow stages) at N = 1024 and 4096."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
import rsp_chains_b200 as R
from rsp_chains_b200.params import CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, MAGParams
torch.cuda.set_device(0)
def wide_params(n, dw, bp, tw, grow=None, **kw):
    s = n.bit_length() - 1
    fp = FFTParams.fixed(dataWidth=dw, twiddleWidth=tw, numPoints=n, binPoint=bp, expandLogic=list(grow or [0] * s), keepMSBorLSB=[True] * s, **kw)
    return FftMagCfarVanillaParameters(fftParams=fp, magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
                                       cfarParams=CFARParams(protoIn=FixedPoint(16, 0), protoThreshold=FixedPoint(16, 0), protoScaler=FixedPoint(16, 0),
                                                             leadLaggWindowSize=16, guardWindowSize=2, fftSize=max(n, 64), CFARAlgorithm=CACFARType))
for n in (1024, 4096):
    for grow in (None, [1] * (n.bit_length() - 1)):
        rows = (1 << 24) // n
        p = wide_params(n, 24, 22, 24, grow, trimType="Convergent", useBitReverse=True)
        with R.FftMagCfarChainVanilla(p) as ch:
            k = ch.fft_out_beat_bytes // 8
            d_in = torch.randint(-(1 << 40), 1 << 40, (rows, n), dtype=torch.int64, device="cuda")
            d_out = torch.zeros((rows, n, k), dtype=torch.int64, device="cuda")
            ms = bench._timed(torch, lambda: ch.fft(d_in, rows, d_out))
            print(json.dumps({"n": n, "grow": bool(grow), "rows": rows, "out_beat_bytes": 8 * k, "ms": round(ms, 4), "Msamples_per_s": round(rows * n / ms / 1e3, 1),
                              "GBps": round(rows * n * (8 + 8 * k) / ms / 1e6, 1)}))
