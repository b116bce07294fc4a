#!/usr/bin/env python
"""Device-resident throughput of every BASELINE.json configuration (C1..C5) on one GPU: Msamples/s, GB/s of algorithmic traffic,
fraction of the measured HBM peak.  `python scripts/bench_configs.py [--quick]` -> one JSON line per configuration.
(bench.py is the contract benchmark; this sweep fills the results table of BASELINE.md.)"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import rsp_chains_b200 as R  # noqa: E402
from rsp_chains_b200 import synth  # noqa: E402
from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, GOSCFARType,  # noqa: E402
                                    MAGParams)


def fixed_params(n, W, G, ctype=CACFARType):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=0),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, 0), protoThreshold=FixedPoint(16, 0), protoScaler=FixedPoint(16, 0),
                              leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype, sendCut=False))


def f32_params(n, W, G):
    return FftMagCfarVanillaParameters(fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True), magParams=MAGParams.fixed(),
                                       cfarParams=CFARParams(leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=CACFARType, sendCut=False))


def timed(fn, steps=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]))


def run_case(name, params, log_n, regs, rows, n_rx, fp32=False, scaler=3.5, distinct=64, bytes_per_sample=8.0, peak=6542.7):
    n = 1 << log_n
    distinct = min(distinct, rows)
    if fp32:
        base = torch.from_numpy(synth.chirps_f32(distinct, n, seed=1)).cuda()
        d_in = base.repeat((rows + distinct - 1) // distinct, 1, 1)[:rows].contiguous()
        d_out = torch.empty((rows, n), dtype=torch.float32, device="cuda")
        d_mask = torch.zeros((rows, n // 32), dtype=torch.int32, device="cuda")
    else:
        base = torch.from_numpy(synth.chirps_fixed(distinct, n, seed=1).view(np.int32)).cuda()
        d_in = base.repeat((rows + distinct - 1) // distinct, 1)[:rows].contiguous()
        d_out = torch.empty_like(d_in)
        d_mask = None
    with R.FftMagCfarChainVanilla(params) as ch:
        ch.configure(log_n, 2, regs)
        if fp32:
            ch.set_threshold_scaler_f32(scaler)
        ms = timed(lambda: ch.run(d_in, rows // n_rx, n_rx, out=d_out, peak_mask=d_mask))
        variant = ch.variant
    samples = rows * n
    gbs = samples * bytes_per_sample / (ms * 1e-3) / 1e9
    print(json.dumps({"config": name, "variant": variant, "rows": rows, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(samples / ms / 1e3, 1),
                      "GBps_algorithmic": round(gbs, 1), "frac_of_measured_hbm_peak": round(gbs / peak, 4), "bytes_per_sample": bytes_per_sample,
                      "footprint_MiB": round(samples * bytes_per_sample / 2 ** 20, 1)}), flush=True)
    del d_in, d_out
    torch.cuda.empty_cache()


def run_generator_case(rows, peak):
    """RspChainVanilla with the PLFG enabled (rsp_run_from_generator: generator kernel -> scratch beats -> fused chain) and the
    PLFG -> NCO generator alone (rsp_generate, 4 B/sample written): a triangle chirp of three segments, RspChainVanillaApp parameters"""
    p = R.rsp_chain_vanilla_app_params()
    n, bb = 1024, p.beatBytes
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    segs = [(1, 200, True, False), (2, 180, False, False), (0, 100, False, True)]   # (slope, nSamples, up, reset)
    W, OW = p.plfgParams.maxNumOfSamplesWidth, p.plfgParams.outputWidthInt

    def ram_word(slope, ns, up, rst):   # {reset, sign, nSamples, slope} (PLFGDspBlockMem.scala:168-178)
        return (int(rst) << (OW + W + 1)) | (int(up) << (OW + W)) | ((ns & ((1 << W) - 1)) << OW) | (slope & ((1 << OW) - 1))

    with R.RspChainVanilla(p) as ch:
        ch.configure(10, 2, regs)
        for i, sg in enumerate(segs):
            ch.memWriteWord(p.plfgRAM.base + i * bb, ram_word(*sg))
        seg0 = 6
        rep0 = seg0 + p.plfgParams.maxChirpOrdinalNum
        ord0 = rep0 + p.plfgParams.maxNumOfDifferentChirps
        for idx, v in ((2, 0), (3, 7), (4, 1), (5, 3), (seg0, len(segs)), (rep0, 2), (ord0, 0), (1, 0), (0, 1)):   # as RspChainVanillaTester.scala:86-130
            ch.memWriteWord(p.plfgAddress.base + idx * bb, v)
        d_out = torch.empty((rows, n), dtype=torch.int32, device="cuda")
        d_nco = torch.empty((rows, n), dtype=torch.int32, device="cuda")
        ms_chain = timed(lambda: ch.run_from_generator(rows, 1, out=d_out))
        ms_gen = timed(lambda: ch.generate(rows * n, d_nco))
        variant = ch.variant
    samples = rows * n
    for name, ms, bps, var in (("PLFG -> NCO -> chain from the generator (C1 parameters), 2 launches", ms_chain, 12.0, "generator + " + variant),
                               ("PLFG -> NCO generator alone", ms_gen, 4.0, "generator_kernel")):
        gbs = samples * bps / (ms * 1e-3) / 1e9
        print(json.dumps({"config": name, "variant": var, "rows": rows, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(samples / ms / 1e3, 1),
                          "GBps_algorithmic": round(gbs, 1), "frac_of_measured_hbm_peak": round(gbs / peak, 4), "bytes_per_sample": bps,
                          "footprint_MiB": round(samples * bps / 2 ** 20, 1)}), flush=True)


def main():
    quick = "--quick" in sys.argv
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    torch.cuda.set_device(0)
    ca = lambda n, W, G, ds, mode=0: dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=ds, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)  # noqa: E731
    run_case("C1 1 x 1024 fixed CA (W32 G4)", fixed_params(1024, 64, 4), 10, ca(1024, 32, 4, 5), 1, 1, peak=peak)
    run_case("C2 one frame 256 x 1024 fixed CA (W16 G4)", fixed_params(1024, 64, 4), 10, ca(1024, 16, 4, 4), 256, 1, peak=peak)
    run_case("C2 x 256 frames (bench.py workload)", fixed_params(1024, 64, 4), 10, ca(1024, 16, 4, 4), 65536, 1, peak=peak)
    for mode, nm in ((1, "GO"), (2, "SO")):
        run_case(f"C3 4 RX x 512 x 2048 fp32 {nm} (W32 G4)", f32_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, mode), 2048, 4, fp32=True,
                 bytes_per_sample=12.125, peak=peak)
    run_case("C3 x 16 frames fp32 GO", f32_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, 1), 32768, 4, fp32=True, bytes_per_sample=12.125, peak=peak)
    run_case("C3 shape in fixed point, GO", fixed_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, 1), 32768, 4, peak=peak)
    os_regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
    run_case("C4 1024 x 4096 fixed GOS (64 cells per side, k=48)", fixed_params(4096, 64, 4, GOSCFARType), 12, os_regs, 1024, 1, peak=peak)
    run_case("C4 x 16: 16384 x 4096 fixed GOS (64 cells per side, k=48)", fixed_params(4096, 64, 4, GOSCFARType), 12, os_regs, 1024 if quick else 16384, 1, peak=peak)
    run_case("C4 shape with CA-CFAR (W64 G4)", fixed_params(4096, 64, 4), 12, ca(4096, 64, 4, 6), 1024 if quick else 16384, 1, peak=peak)
    run_case("C5 8 RX x 4096 x 16384 fixed CA (W64 G8), one frame", fixed_params(16384, 64, 8), 14, ca(16384, 64, 8, 6),
             4096 if quick else 32768, 8, peak=peak)
    run_case("C5 shape fp32 CA, 1/4 frame", f32_params(16384, 64, 8), 14, ca(16384, 64, 8, 6), 8192, 8, fp32=True, bytes_per_sample=12.125, peak=peak)
    run_generator_case(4096 if quick else 65536, peak)


if __name__ == "__main__":
    main()
