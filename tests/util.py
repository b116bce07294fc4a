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


def random_fp32_configuration(rng, trial, sizes=(6, 8, 10)):
    """one random configuration of the DspReal flavour over the four CFAR cores: (params, log_n, cfar registers, scaler, float32 beats [rows, n, 2],
    description) or None when the draw has no valid window.  Shared by the CPU (emulated) and GPU randomised tests."""
    from rsp_chains_b200 import synth
    from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, GOSCFARType, MAGParams, OSGoldenCFARType)
    core = ["mem", "gos", "cash", "os"][int(rng.integers(0, 4))]
    log_n = int(rng.choice(list(sizes)))
    n = 1 << log_n
    W = 16 if core == "cash" else (1 << int(rng.integers(2, 5)) if core == "mem" else int(rng.integers(4, 17)))
    G = int(rng.integers(1, 4))
    lin = int(rng.integers(0, 2))
    kl, kr = int(rng.integers(1, W + 1)), int(rng.integers(1, W + 1))
    pg = int(rng.random() < 0.3) if core != "os" else 0
    mode = 3 if core == "cash" else int(rng.integers(0, 3))
    scaler = float(np.float32(rng.uniform(2.0, 5.0) if lin else rng.uniform(0.0, 2e-4)))
    noise = 2.0 ** -int(rng.integers(3, 7))
    if n <= 2 * W + 2 * G + 1:
        return None
    ctype = {"mem": CACFARType, "gos": GOSCFARType, "cash": CACFARType, "os": OSGoldenCFARType}[core]
    p = FftMagCfarVanillaParameters(fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True), magParams=MAGParams.fixed(),
                                    cfarParams=CFARParams(leadLaggWindowSize=16, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype, sendCut=False,
                                                          includeCASH=core == "cash", minSubWindowSize=4 if core == "cash" else None))
    regs = dict(fftWin=n, logOrLinearMode=lin, divSum=2 if core == "cash" else int(np.log2(W)), peakGrouping=pg, cfarMode=mode, windowCells=W, guardCells=G,
                indexLagg=kl, indexLead=kr, subWindowSize=4)
    xc = synth.chirps_complex(3, n, seed=900 + trial, noise=noise)
    x32 = np.ascontiguousarray(np.stack([xc.real, xc.imag], axis=-1).astype(np.float32))
    return p, log_n, regs, scaler, x32, (trial, core, dict(log_n=log_n, W=W, G=G, lin=lin, pg=pg, kl=kl, kr=kr, scaler=scaler), regs)


def check_fp32_against_f64(O, what, log_n, regs, scaler, x32, thr, pk, tol=1e-4):
    """thresholds within tol (relative) of the fp64 dataflow; detections identical wherever the decision is not within 10 tol of a tie"""
    from oracle.glue import f64_chain
    core, W, G = what[1], regs["windowCells"], regs["guardCells"]
    kw = {"mem": {}, "gos": dict(core=O.CORE_LIS, index_lagg=regs["indexLagg"], index_lead=regs["indexLead"]), "cash": dict(core=O.CORE_ASR, sub_cells=4),
          "os": dict(core=O.CORE_OS)}[core]
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    mag, want_thr, want_pk = f64_chain(O, xq, log_n, 2, regs, float(scaler), W, G, **kw)
    # denominator: the threshold itself, floored at 1e-3 of the largest threshold and at 1e-2 of the row's largest magnitude -- the fp32 transform's own
    # rounding noise is about eps * log2(N) * peak = 1e-6 of the row's peak, whatever the bin holds, so a threshold far below the peak (a small ordered
    # statistic of the noise floor next to a strong target) cannot be held to 1e-4 of its own value in fp32
    floor = np.maximum(1e-3 * np.abs(want_thr).max(), 1e-2 * np.abs(mag).max(axis=1, keepdims=True))
    rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), floor)
    assert rel.max() < tol, (what, float(rel.max()))
    sure = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), floor) > 10 * tol
    if regs.get("peakGrouping"):
        nb = np.minimum(np.abs(mag - np.roll(mag, -1, axis=1)), np.abs(mag - np.roll(mag, -G, axis=1))) / np.maximum(np.abs(mag), floor) > 10 * tol
        sure &= nb
    assert (pk[sure] == want_pk[sure]).all(), what
