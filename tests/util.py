"""Glue between the product's host-side parameter classes and the CPU oracle's configs (tests only)."""
from __future__ import annotations

import numpy as np

from rsp_chains_b200.params import log2Up

from oracle.glue import oracle_cfgs  # noqa: F401


def configure(chain, log_n, sel, cfar_regs: dict):
    """write the run-time registers through the bus exactly as the reference testers do"""
    chain.configure(log_n, sel, cfar_regs)


def to_dev(torch, a: np.ndarray):
    if a.dtype == np.uint32:
        return torch.from_numpy(a.view(np.int32)).cuda()
    return torch.from_numpy(a).cuda()


def from_dev_u32(t) -> np.ndarray:
    return t.cpu().numpy().view(np.uint32)


def decode_cfar_words(words, fft_size):
    """threshold = out >> (log2 N + 1), bin, peak = out & 1 (src/test/scala/FftMagCfarChainTester.scala:163-167)"""
    w = np.asarray(words, dtype=np.uint32)
    b = log2Up(fft_size)
    return (w >> (b + 1)), (w >> 1) & ((1 << b) - 1), w & 1


NCO_FLAVOURS = [dict(), dict(phaseAccEnable=False), dict(phaseWidth=14, tableSize=1024), dict(phaseWidth=14, tableSize=256, tableWidth=12),
                dict(phaseWidth=12, tableSize=256), dict(rasterizedMode=True, tableSize=100), dict(rasterizedMode=True, tableSize=384)]


def random_plfg_programme(rng):
    """one random PLFG programme within RspChainVanillaApp's PLFGParams: (nco keywords, chirps, order, repeats, frameNum, pause, startingPoint,
    samples to ask of an endless schedule).  Shared by the CPU (emulated) and GPU randomised generator tests."""
    nco_kw = NCO_FLAVOURS[int(rng.integers(0, len(NCO_FLAVOURS)))]
    n_types = int(rng.integers(1, 5))
    chirps = [[(int(rng.integers(0, 1 << int(rng.integers(1, 12)))), int(rng.integers(0, 256)), bool(rng.integers(0, 2)), bool(rng.random() < 0.15))
               for _ in range(int(rng.integers(1, 5)))] for _ in range(n_types)]
    n_order = int(rng.integers(1, 9))
    order = [int(rng.integers(0, n_types)) for _ in range(n_order)]
    repeats = [int(rng.integers(0, 5)) for _ in range(n_order)]
    frame_num, inter, start = int(rng.integers(0, 4)), int(rng.integers(0, 64)), int(rng.integers(0, 1 << 16))
    return nco_kw, chirps, order, repeats, frame_num, inter, start, int(rng.integers(3000, 40000))


TRIM_TYPES = ["RoundHalfUp", "RoundDown", "Convergent", "RoundUp", "RoundTowardsZero", "RoundTowardsInfinity", "RoundHalfDown", "RoundHalfTowardsZero",
              "RoundHalfTowardsInfinity", "RoundHalfToOdd"]


def random_generic_configuration(rng, trial, fft_dir_reg=True, max_log=9):
    """one random configuration of the fixed-point chain over its whole parameter space (what only the run-time-parameter kernel serves included):
    (params, log_n, sel, cfar registers, beats, fftDir, description) or None when the draw has no valid window.  Shared by the CPU (emulated) and
    GPU randomised tests."""
    import rsp_chains_b200 as R
    from rsp_chains_b200 import synth
    from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, GOSCACFARType, GOSCFARType,
                                        MAGParams, OSGoldenCFARType)
    log_np = int(rng.integers(3, max_log + 1))
    radix, dit = ("2^2", False) if rng.random() < 0.5 else (str(rng.choice(["2", "2^2"])), bool(rng.integers(0, 2)))
    # a run-time size below numPoints (radix 2^2 drops an even number of stages, Radix22RunTimeSpec.scala:18-19)
    log_n = log_np - 2 * int(rng.integers(0, (log_np - 3) // 2 + 1)) if (radix == "2^2" and not dit and rng.random() < 0.4) else log_np
    n, npts = 1 << log_n, 1 << log_np
    bp = int(rng.choice([0, 0, 3, 8, 12, 14]))
    trim = str(rng.choice(TRIM_TYPES))
    use4 = bool(rng.integers(0, 2))
    keep = [bool(rng.random() < 0.8) for _ in range(log_np)]
    ctype = [CACFARType, GOSCFARType, GOSCACFARType, OSGoldenCFARType][int(rng.integers(0, 4))]
    cash = bool(ctype == CACFARType and rng.random() < 0.3)
    w_max = max(2, min(32, (n - 4) // 4))
    W = 16 if cash else int(rng.integers(2, w_max + 1))
    if ctype == CACFARType and not cash:
        W = 1 << int(np.log2(W))
    G = int(rng.integers(1, 4))
    sel = int(rng.integers(0, 4))
    lin = int(rng.integers(0, 2)) if sel != 3 else 0
    pg = int(rng.random() < 0.3) if ctype != OSGoldenCFARType else 0
    send_cut = bool(rng.random() < 0.15)
    fft_dir = int(rng.random() < 0.85)
    gen_g = 3 if rng.random() < 0.5 else G
    scaler = int(rng.integers(0, 40))
    cfar_mode, kl, kr = int(rng.integers(0, 3)), int(rng.integers(1, W + 1)), int(rng.integers(1, W + 1))
    if n <= 2 * W + 2 * G + 1 or (cash and n < 64):
        return None
    gen_w = max(4, 1 << int(np.ceil(np.log2(W))))      # the generator's window: a power of two above the guard's
    p = FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=npts, useBitReverse=True, runTime=True, binPoint=bp, trimType=trim, use4Muls=use4,
                                  sdfRadix=radix, decimType=R.DITDecimType if dit else R.DIFDecimType, keepMSBorLSB=keep,
                                  fftDirReg=bool(fft_dir_reg), fftDir=True if fft_dir_reg else bool(fft_dir)),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=bp, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, bp), protoThreshold=FixedPoint(16, bp), protoScaler=FixedPoint(16, bp), leadLaggWindowSize=gen_w,
                              guardWindowSize=gen_g, fftSize=npts, includeCASH=cash, minSubWindowSize=4 if cash else None, CFARAlgorithm=ctype,
                              sendCut=send_cut))
    regs = dict(fftWin=n, thresholdScaler=int(3.0 * (1 << bp)) if lin else scaler, logOrLinearMode=lin, divSum=2 if cash else max(0, int(np.log2(W))),
                peakGrouping=pg, cfarMode=cfar_mode, cfarAlgorithm=0, windowCells=W, guardCells=G, indexLagg=kl, indexLead=kr, subWindowSize=4 if cash else W)
    beats = np.concatenate([synth.chirps_fixed(2, n, seed=500 + trial, noise=2.0 ** -4), synth.random_beats(2, n, seed=600 + trial)])
    what = (trial, dict(npts=npts, n=n, radix=radix, dit=dit, bp=bp, trim=trim, use4=use4, keep=keep, ctype=ctype, cash=cash, W=W, G=G, gen=(gen_w, gen_g),
                        sel=sel, lin=lin, pg=pg, send_cut=send_cut, fft_dir=fft_dir), regs)
    return p, log_n, sel, regs, beats, fft_dir, what
