"""Maps the product's host-side parameter classes + run-time registers onto the oracle's configs (TEST INFRASTRUCTURE:
used by tests/, __graft_entry__.smoke() and bench.py's CPU legs only)."""
from __future__ import annotations

from rsp_chains_b200.params import CACFARType, DITDecimType, GOSCFARType, OSGoldenCFARType, log2Up


def oracle_cfgs(O, params, log_n, sel, cfar_regs: dict):
    f, m, c = params.fftParams, params.magParams, params.cfarParams
    smax = log2Up(f.numPoints)
    mask = 0
    for i, k in enumerate(f.keepMSBorLSB):
        mask |= (1 if k else 0) << i
    dit = f.decimType == DITDecimType
    # a run-time size of k stages is the LAST k stages in DIF, the FIRST k in DIT (SDFChainRadix2.scala:127-134,196-203)
    mask = (mask if dit else (mask >> (smax - log_n))) & ((1 << log_n) - 1)
    # grow stages (DIF): a run-time size keeps the last log_n stages; the ones in front have already widened the types
    emask = sum((1 if e else 0) << i for i, e in enumerate(f.expandLogic))
    expand = (emask >> (smax - log_n)) & ((1 << log_n) - 1)
    base_growth = bin(emask & ((1 << (smax - log_n)) - 1)).count("1")
    fft = O.fft_cfg(log_n, f.dataWidth, f.binPoint, f.twiddleWidth, f.trimType, f.use4Muls, f.useBitReverse, mask, f.fftDir,
                    f.sdfRadix, dit, smax, expand_mask=expand, base_growth=base_growth)
    mag = O.mag_cfg(m.dataWidth, m.binPoint, m.dataWidthLog, m.binPointLog, m.log2LookUpWidth, sel)
    if c.CFARAlgorithm == CACFARType:
        core = O.CORE_ASR if c.includeCASH else O.CORE_MEM
        alg = 0
    elif c.CFARAlgorithm == OSGoldenCFARType:
        core, alg = O.CORE_OS, 0
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


# ---------------------------------------------------------------------------------------------------------------------
# fp64 dataflow of the DspReal flavour and small helpers shared by tests/ and __graft_entry__.smoke()
def f64_chain(O, x, log_n, sel, regs, scaler, W, G, core=None, index_lagg=1, index_lead=1, sub_cells=None, guard_window_size=None):
    """FFT(x)/N in natural order, magnitude (sel 1 |z|^2, 2 JPL, 3 log2 of JPL), then the CFAR core's threshold / peak in doubles
    with the hardware's edge policy: CFARCoreWithMem one-sided windows, CFARCoreWithLis / ASR zero unless both windows are full
    (CFARCoreWithLis.scala:172-173).  x: complex128 [rows, n].  Returns (mag, thr, peak)."""
    import numpy as np
    n = 1 << log_n
    core = O.CORE_MEM if core is None else core
    thr = np.zeros((x.shape[0], n)); peak = np.zeros((x.shape[0], n), np.uint8); mag = np.zeros((x.shape[0], n))
    cfg = O.cfar_cfg(n, W, G, 0, core=core, cfar_mode=regs["cfarMode"], div_sum=regs.get("divSum", 0),
                     log_or_linear=regs["logOrLinearMode"], peak_grouping=regs.get("peakGrouping", 0),
                     cfar_algorithm=1 if core == O.CORE_LIS and regs.get("cfarAlgorithm", 1) else 0,
                     index_lagg=index_lagg, index_lead=index_lead, sub_cells=sub_cells,
                     guard_window_size=guard_window_size if guard_window_size is not None else G)
    for r in range(x.shape[0]):
        fr, fi = O.fft_f64(log_n, x[r].real, x[r].imag)
        if sel == 2:
            m = np.array([O.jpl_f64(a, b) for a, b in zip(fr, fi)])
        elif sel == 1:
            m = fr * fr + fi * fi
        else:
            m = np.log2(np.array([O.jpl_f64(a, b) for a, b in zip(fr, fi)]))
        mag[r] = m
        thr[r], peak[r] = O.cfar_f64(cfg, scaler, m)
    return mag, thr, peak


def unpack_mask(mask, n):
    """peak-mask words [rows, n/32] -> bits [rows, n]"""
    import numpy as np
    bits = np.unpackbits(mask.view(np.uint8), bitorder="little").reshape(mask.shape[0], -1)
    return bits[:, :n]


def write_plfg(ch, p, O, starting_point, chirps, order=None, repeats=None, frame_num=1, inter=10):
    """program the PLFG through the register bus as the reference tester does (RspChainVanillaTester.scala:86-130).
    chirps: list (by chirp ordinal number) of segment lists [(slope, nSamples, up, reset), ...]; order: chirpOrdinalNums per
    different chirp (default 0, 1, ...); repeats: repeatedChirpNums per different chirp (default 1).  Returns the RAM image."""
    g = p.plfgParams
    bb = p.beatBytes
    order = list(range(len(chirps))) if order is None else list(order)
    repeats = [1] * len(order) if repeats is None else list(repeats)
    ram = [0] * (g.maxChirpOrdinalNum * g.maxNumOfSegments)
    for c, segs in enumerate(chirps):
        for i, (slope, ns, up, rst) in enumerate(segs):
            ram[c * g.maxNumOfSegments + i] = O.plfg_ram_word(slope, ns, up, rst, output_width=g.outputWidthInt, samples_width=g.maxNumOfSamplesWidth)
            ch.memWriteWord(p.plfgRAM.base + (c * g.maxNumOfSegments + i) * bb, ram[c * g.maxNumOfSegments + i])
    seg0 = 6
    rep0 = seg0 + g.maxChirpOrdinalNum
    ord0 = rep0 + g.maxNumOfDifferentChirps
    ch.memWriteWord(p.plfgAddress.base + 2 * bb, frame_num)
    ch.memWriteWord(p.plfgAddress.base + 3 * bb, inter)
    ch.memWriteWord(p.plfgAddress.base + 4 * bb, len(order))
    ch.memWriteWord(p.plfgAddress.base + 5 * bb, starting_point)
    for c, segs in enumerate(chirps):
        ch.memWriteWord(p.plfgAddress.base + (seg0 + c) * bb, len(segs))
    for d, (o, r) in enumerate(zip(order, repeats)):
        ch.memWriteWord(p.plfgAddress.base + (rep0 + d) * bb, r)
        ch.memWriteWord(p.plfgAddress.base + (ord0 + d) * bb, o)
    ch.memWriteWord(p.plfgAddress.base + 1 * bb, 0)
    ch.memWriteWord(p.plfgAddress.base + 0 * bb, 1)   # enable
    return ram


def oracle_plfg(O, p, starting_point, chirps, order=None, repeats=None, frame_num=1, inter=10):
    """the oracle PLFG (RTL pause model) for the same programme as write_plfg"""
    g = p.plfgParams
    order = list(range(len(chirps))) if order is None else list(order)
    repeats = [1] * len(order) if repeats is None else list(repeats)
    ram = [0] * (g.maxChirpOrdinalNum * g.maxNumOfSegments)
    for c, segs in enumerate(chirps):
        for i, sg in enumerate(segs):
            ram[c * g.maxNumOfSegments + i] = O.plfg_ram_word(*sg, output_width=g.outputWidthInt, samples_width=g.maxNumOfSamplesWidth)
    return O.Plfg(max_num_of_segments=g.maxNumOfSegments, max_num_of_samples_width=g.maxNumOfSamplesWidth,
                  output_width=g.outputWidthInt + g.outputWidthFrac, output_width_int=g.outputWidthInt,
                  max_chirp_ordinal_num=g.maxChirpOrdinalNum, max_num_of_different_chirps=g.maxNumOfDifferentChirps,
                  frame_num=frame_num, interframe_num_of_samples=inter, different_chirps_num=len(order), starting_point=starting_point,
                  segment_nums=[len(s) for s in chirps], repeated_chirp_nums=repeats, chirp_ordinal_nums=order, ram=ram)
