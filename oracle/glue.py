"""Maps the product's host-side parameter classes + run-time registers onto the oracle's configs (TEST INFRASTRUCTURE:
used by tests/, __graft_entry__.smoke() and bench.py's CPU legs only)."""
from __future__ import annotations

from rsp_chains_b200.params import CACFARType, DITDecimType, GOSCFARType, log2Up


def oracle_cfgs(O, params, log_n, sel, cfar_regs: dict):
    f, m, c = params.fftParams, params.magParams, params.cfarParams
    smax = log2Up(f.numPoints)
    mask = 0
    for i, k in enumerate(f.keepMSBorLSB):
        mask |= (1 if k else 0) << i
    dit = f.decimType == DITDecimType
    # a run-time size of k stages is the LAST k stages in DIF, the FIRST k in DIT (SDFChainRadix2.scala:127-134,196-203)
    mask = (mask if dit else (mask >> (smax - log_n))) & ((1 << log_n) - 1)
    fft = O.fft_cfg(log_n, f.dataWidth, f.binPoint, f.twiddleWidth, f.trimType, f.use4Muls, f.useBitReverse, mask, f.fftDir,
                    f.sdfRadix, dit, smax)
    mag = O.mag_cfg(m.dataWidth, m.binPoint, m.dataWidthLog, m.binPointLog, m.log2LookUpWidth, sel)
    if c.CFARAlgorithm == CACFARType:
        core = O.CORE_ASR if c.includeCASH else O.CORE_MEM
        alg = 0
    else:
        core = O.CORE_LIS
        alg = 1 if c.CFARAlgorithm == GOSCFARType else cfar_regs.get("cfarAlgorithm", 0)
    r = dict(fftWin=c.fftSize, thresholdScaler=0, logOrLinearMode=0, divSum=0, peakGrouping=0, cfarMode=0,
             windowCells=c.leadLaggWindowSize, guardCells=c.guardWindowSize, indexLagg=c.leadLaggWindowSize // 2,
             indexLead=c.leadLaggWindowSize // 2, subWindowSize=c.leadLaggWindowSize)
    r.update(cfar_regs)
    cfar = O.cfar_cfg(r["fftWin"], r["windowCells"], r["guardCells"], r["thresholdScaler"], core=core,
                      in_width=c.protoIn.width, in_bp=c.protoIn.binaryPoint, thr_width=c.protoThreshold.width,
                      thr_bp=c.protoThreshold.binaryPoint, scaler_width=c.protoScaler.width, scaler_bp=c.protoScaler.binaryPoint,
                      lead_lagg_window_size=c.leadLaggWindowSize, guard_window_size=c.guardWindowSize, fft_size=c.fftSize,
                      min_sub_window_size=c.minSubWindowSize or 0, send_cut=c.sendCut, log_or_linear=r["logOrLinearMode"],
                      div_sum=r["divSum"], peak_grouping=r["peakGrouping"], cfar_algorithm=alg, cfar_mode=r["cfarMode"],
                      index_lagg=r["indexLagg"], index_lead=r["indexLead"], sub_cells=r["subWindowSize"])
    return fft, mag, cfar


def apply_window(O, params, beats):
    """the window multiply in front of the forward transform (identity for WindowFunctionTypes.None / the IFFT)"""
    f = params.fftParams
    w = f.windowFunc
    if w is None or w.kind == 0 or not f.fftDir:
        return beats
    raw = O.window_raw(O.window_coeffs(w.kind, f.numPoints, w.a0, w.a1, w.a2, w.dv), w.dataWidth)
    return O.apply_window_words(beats, raw, w.dataWidth, f.dataWidth)
