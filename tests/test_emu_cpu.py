"""CPU emulation of the run-time-parameter kernels (tests/emu/emu.cpp) against the oracle, without a GPU.

The per-thread phases of rsp_device*.cuh are compiled with g++ and run thread by thread in the kernels' phase order, on argument
blocks built by the product's own planner (rsp_host.hpp).  This holds the index maps, the integer arithmetic and the host planning of
every configuration the run-time-parameter kernels serve to the oracle on the GPU-less box (the `-m gpu` tests then only add the CUDA
launch and the specialised kernels).  The harness is test infrastructure: the product library never contains it."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import _abi, synth
from rsp_chains_b200.chain import _c_params
from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, GOSCACFARType, GOSCFARType,
                                    MAGParams, OSGoldenCFARType)
from oracle.glue import f64_chain, oracle_cfgs, oracle_plfg, unpack_mask

HERE = os.path.dirname(os.path.abspath(__file__))
STAGE_FFT, STAGE_MAG, STAGE_CFAR = 0, 1, 2
CFAR_REGS = ["fftWin", "thresholdScaler", "logOrLinearMode", "divSum", "peakGrouping", "cfarAlgorithm", "cfarMode", "windowCells", "guardCells",
             "indexLagg", "indexLead", "subWindowSize"]


@pytest.fixture(scope="session")
def emu():
    subprocess.run(["make", "-C", os.path.join(HERE, "emu")], check=True, capture_output=True)
    L = C.CDLL(os.path.join(HERE, "emu", "librsp_emu.so"))
    L.rsp_emu_run_fixed.restype = C.c_int
    L.rsp_emu_run_f32.restype = C.c_int
    L.rsp_emu_generate.restype = C.c_int
    return L


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def cfar_reg_words(p, regs):
    c = p.cfarParams
    r = dict(fftWin=c.fftSize, thresholdScaler=0, logOrLinearMode=0, divSum=0, peakGrouping=0, cfarAlgorithm=0, cfarMode=0,
             windowCells=c.leadLaggWindowSize, guardCells=c.guardWindowSize, indexLagg=c.leadLaggWindowSize // 2,
             indexLead=c.leadLaggWindowSize // 2, subWindowSize=c.leadLaggWindowSize)
    r.update(regs)
    return _u32([r[k] & 0xFFFFFFFF for k in CFAR_REGS])


def emu_fixed(emu, p, log_n, sel, regs, beats, first=STAGE_FFT, last=STAGE_CFAR, fft_dir=1, keep=0, dowin=0, win_ram=None, want_dets=False):
    cp = _c_params(p.fftParams, p.magParams, p.cfarParams)
    beats = _u32(beats)
    out = np.zeros_like(beats)
    fft_regs = _u32([log_n, fft_dir, keep, dowin])
    dets = (_abi.Detection * 65536)()
    cnt = C.c_uint32(0)
    ovf = C.c_uint32(0)
    mode = C.c_int(-1)
    err = C.create_string_buffer(512)
    wr = np.ascontiguousarray(win_ram, dtype=np.int32) if win_ram is not None else None
    rc = emu.rsp_emu_run_fixed(C.byref(cp), fft_regs.ctypes.data_as(C.c_void_p), _u32([sel]).ctypes.data_as(C.c_void_p),
                               cfar_reg_words(p, regs).ctypes.data_as(C.c_void_p), first, last, beats.ctypes.data_as(C.c_void_p),
                               C.c_uint32(beats.shape[0]), out.ctypes.data_as(C.c_void_p), dets if want_dets else None,
                               C.byref(cnt) if want_dets else None, C.c_uint32(65536), wr.ctypes.data_as(C.c_void_p) if wr is not None else None,
                               C.byref(ovf), C.byref(mode), err, 512)
    if rc:
        raise RuntimeError(f"emu rc={rc}: {err.value.decode()}")
    d = np.frombuffer(dets, dtype=np.int32).reshape(-1, 4)[:cnt.value].copy() if want_dets else None
    return out, ovf.value, mode.value, d


def chain_params(n, bp=0, trim="RoundHalfUp", ctype=CACFARType, cash=False, W=16, G=2, send_cut=False, **fft_kw):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=bp, trimType=trim, **fft_kw),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=bp, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, bp), protoThreshold=FixedPoint(16, bp), protoScaler=FixedPoint(16, bp), leadLaggWindowSize=W,
                              guardWindowSize=G, fftSize=n, includeCASH=cash, minSubWindowSize=4 if cash else None, CFARAlgorithm=ctype, sendCut=send_cut))


@pytest.mark.parametrize("log_n", [1, 2, 3, 4, 5, 6, 8, 10])
def test_emu_chain_all_sizes_bp0(emu, oracle, log_n):
    """binPoint 0 / RoundHalfUp (the chain apps' configuration: kernel mode FAST on the run-time-parameter kernel), full-range random beats"""
    n = 1 << log_n
    W = max(1, min(16, n // 8)); G = 1
    p = chain_params(n, 0, W=max(2, W), G=1)
    beats = synth.random_beats(9, n, seed=log_n)
    if not (n > 2 * W + 2 * G + 1):
        fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, dict(fftWin=n))
        got, _, _, _ = emu_fixed(emu, p, log_n, 2, dict(fftWin=n), beats, last=STAGE_FFT)
        assert (got == oracle.fft_fixed_words(fft, beats)).all()
        return
    for mode in (0, 1, 2):
        regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=max(0, W.bit_length() - 1), peakGrouping=mode & 1, cfarMode=mode,
                    windowCells=W, guardCells=G)
        got, _, kmode, _ = emu_fixed(emu, p, log_n, 2, regs, beats)
        fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
        assert kmode == 1      # MODE_FAST
        assert (got == oracle.chain_fixed_words(fft, mag, cfar, beats)).all(), f"mode {mode}"


@pytest.mark.parametrize("bp,trim,use4,sel", [(12, "RoundHalfUp", False, 2), (14, "Convergent", False, 1), (12, "RoundDown", True, 3), (14, "RoundHalfToOdd", False, 2),
                                              (12, "RoundTowardsZero", True, 0), (0, "Convergent", False, 2)])
def test_emu_binpoints_trims_magnitudes(emu, oracle, bp, trim, use4, sel):
    """kernel modes FAST3 / GENERIC: 3- and 4-multiplier products, every magnitude mode, log threshold"""
    p = chain_params(256, bp, trim, use4Muls=use4)
    regs = dict(fftWin=256, thresholdScaler=700, logOrLinearMode=0, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=2)
    beats = np.concatenate([synth.chirps_fixed(5, 256, seed=3), synth.random_beats(4, 256, seed=4)])
    got, _, _, _ = emu_fixed(emu, p, 8, sel, regs, beats)
    fft, mag, cfar = oracle_cfgs(oracle, p, 8, sel, regs)
    assert (got == oracle.chain_fixed_words(fft, mag, cfar, beats)).all()


@pytest.mark.parametrize("ctype,cash,mode", [(GOSCFARType, False, 0), (GOSCACFARType, False, 1), (CACFARType, True, 3), (CACFARType, True, 2), (OSGoldenCFARType, False, 0)])
@pytest.mark.parametrize("send_cut", [False, True])
def test_emu_cfar_cores(emu, oracle, ctype, cash, mode, send_cut):
    """Lis (GOS and its CA mode), ASR (CASH and sub-window averages) and the combined-window OS statistic, detection list, sendCut"""
    p = chain_params(1024, 0, ctype=ctype, cash=cash, W=16, G=2, send_cut=send_cut)
    regs = dict(fftWin=1024, thresholdScaler=3, logOrLinearMode=1, divSum=2 if cash else 4, peakGrouping=1 if ctype != OSGoldenCFARType else 0,
                cfarMode=mode, cfarAlgorithm=0, windowCells=16, guardCells=2, indexLagg=12, indexLead=5, subWindowSize=4)
    beats = np.concatenate([synth.chirps_fixed(3, 1024, seed=8), synth.random_beats(2, 1024, seed=9)])
    got, _, _, dets = emu_fixed(emu, p, 10, 2, regs, beats, want_dets=True)
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want, _, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    assert (got == want).all()
    if not send_cut:
        exp = set(map(tuple, np.argwhere((want & 1) == 1).tolist()))
        assert set((int(d[0]), int(d[1])) for d in dets) == exp and len(dets) == len(exp)


@pytest.mark.parametrize("radix,dit", [("2", False), ("2", True), ("2^2", True)])
@pytest.mark.parametrize("log_n", [3, 6, 10])
def test_emu_radix2_and_dit(emu, oracle, radix, dit, log_n):
    n = 1 << log_n
    p = chain_params(n, 12, "Convergent", sdfRadix=radix, decimType=R.DITDecimType if dit else R.DIFDecimType, W=max(2, n // 16), G=1)
    beats = synth.random_beats(6, n, seed=log_n + 50)
    fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, dict(fftWin=n))
    got, _, kmode, _ = emu_fixed(emu, p, log_n, 2, dict(fftWin=n), beats, last=STAGE_FFT)
    assert kmode == 3      # MODE_RADIX2
    assert (got == oracle.fft_fixed_words(fft, beats)).all()


def test_emu_runtime_size_keep_lsb_ifft_window_overflow(emu, oracle):
    """a run-time size below numPoints, keep-LSB stages through the register, IFFT, a window ROM, the window RAM of FFTBlockWithWindowing,
    and the overflow status word"""
    from rsp_chains_b200 import WindowFunctionTypes as WT
    beats = np.concatenate([synth.chirps_fixed(4, 256, seed=12), synth.random_beats(3, 256, seed=13)])
    # run-time size 256 of 1024 with keepMSBorLSBReg: physical stages 2..9
    p = chain_params(1024, 0, keepMSBorLSBReg=True, overflowReg=True)
    keep = 0b1101111011
    got, ovf, _, _ = emu_fixed(emu, p, 8, 2, dict(fftWin=256), beats, last=STAGE_FFT, keep=keep)
    p.fftParams.keepMSBorLSB = [bool((keep >> i) & 1) for i in range(10)]
    fft, _, _ = oracle_cfgs(oracle, p, 8, 2, dict(fftWin=256))
    want, want_ovf = oracle.fft_fixed_words_ovf(fft, beats)
    assert (got == want).all() and ovf == want_ovf << 2 and want_ovf != 0
    # IFFT through the fftDir register
    p = chain_params(256, 12, "Convergent", fftDirReg=True)
    got, _, _, _ = emu_fixed(emu, p, 8, 2, dict(fftWin=256), beats, last=STAGE_FFT, fft_dir=0)
    p.fftParams.fftDir = False
    fft, _, _ = oracle_cfgs(oracle, p, 8, 2, dict(fftWin=256))
    assert (got == oracle.fft_fixed_words(fft, beats)).all()
    # window ROM
    p = chain_params(256, 0, windowFunc=WT.Hamming(16))
    got, _, _, _ = emu_fixed(emu, p, 8, 2, dict(fftWin=256), beats, last=STAGE_FFT)
    raw = oracle.window_raw(oracle.window_coeffs(1, 256, 0.54, 0.46), 16)
    fft, _, _ = oracle_cfgs(oracle, p, 8, 2, dict(fftWin=256))
    assert (got == oracle.fft_fixed_words(fft, oracle.apply_window_words(beats, raw, 16, 16))).all()
    # window RAM (FFTBlockWithWindowing), enabled and disabled
    p = chain_params(256, 0)
    p.fftParams.windowRam = True
    ram = [int(c * (1 << 14)) for c in oracle.window_coeffs(3, 256, 0.42, 0.5, 0.08)]
    got, _, _, _ = emu_fixed(emu, p, 8, 2, dict(fftWin=256), beats, last=STAGE_FFT, dowin=1, win_ram=ram)
    assert (got == oracle.fft_fixed_words(fft, oracle.apply_window_words(beats, np.array(ram), 16, 16))).all()
    got, _, _, _ = emu_fixed(emu, p, 8, 2, dict(fftWin=256), beats, last=STAGE_FFT, dowin=0, win_ram=ram)
    assert (got == oracle.fft_fixed_words(fft, beats)).all()


def test_emu_blocks_alone(emu, oracle):
    """LogMagMux and CFAR as single blocks (first / last stage selection of the kernel)"""
    p = chain_params(256, 0)
    regs = dict(fftWin=256, thresholdScaler=3, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=2)
    prof = np.floor(np.random.default_rng(1).exponential(200.0, (4, 256))).astype(np.uint32)
    got, _, _, _ = emu_fixed(emu, p, 8, 2, regs, prof, first=STAGE_CFAR, last=STAGE_CFAR)
    _, _, cfar = oracle_cfgs(oracle, p, 8, 2, regs)
    assert (got == oracle.cfar_words(cfar, prof)).all()


@pytest.mark.parametrize("core", ["mem", "gos", "cash", "os"])
def test_emu_fp32_flavour(emu, oracle, core):
    """the DspReal flavour's phases in float on the CPU against the fp64 dataflow: 1e-4 relative, identical detections"""
    log_n, W, G = 8, 16, 2
    n = 1 << log_n
    ctype = {"mem": CACFARType, "gos": GOSCFARType, "cash": CACFARType, "os": OSGoldenCFARType}[core]
    p = FftMagCfarVanillaParameters(fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True), magParams=MAGParams.fixed(),
                                    cfarParams=CFARParams(leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype, sendCut=False,
                                                          includeCASH=core == "cash", minSubWindowSize=4 if core == "cash" else None))
    regs = dict(fftWin=n, logOrLinearMode=1, divSum=2 if core == "cash" else 4, peakGrouping=0, cfarMode=3 if core == "cash" else 1, windowCells=W, guardCells=G,
                indexLagg=12, indexLead=6, subWindowSize=4)
    xc = synth.chirps_complex(5, n, seed=4)
    x32 = np.ascontiguousarray(np.stack([xc.real, xc.imag], axis=-1).astype(np.float32))
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    cp = _c_params(p.fftParams, p.magParams, p.cfarParams)
    thr = np.zeros((5, n), np.float32); mask = np.zeros((5, n // 32), np.uint32)
    err = C.create_string_buffer(512)
    rc = emu.rsp_emu_run_f32(C.byref(cp), _u32([log_n, 1, 0, 0]).ctypes.data_as(C.c_void_p), _u32([2]).ctypes.data_as(C.c_void_p),
                             cfar_reg_words(p, regs).ctypes.data_as(C.c_void_p), C.c_float(3.0), STAGE_FFT, STAGE_CFAR, x32.ctypes.data_as(C.c_void_p),
                             C.c_uint32(5), thr.ctypes.data_as(C.c_void_p), mask.ctypes.data_as(C.c_void_p), err, 512)
    assert rc == 0, err.value.decode()
    kw = {"mem": {}, "gos": dict(core=oracle.CORE_LIS, index_lagg=12, index_lead=6), "cash": dict(core=oracle.CORE_ASR, sub_cells=4), "os": dict(core=oracle.CORE_OS)}[core]
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, 3.0, W, G, **kw)
    scale = np.abs(want_thr).max()
    assert (np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale)).max() < 1e-4
    pk = unpack_mask(mask, n)
    sure = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale) > 1e-3
    assert (pk[sure] == want_pk[sure]).all() and want_pk.sum() >= 5


def emu_generate(emu, p, oracle, chirps, order, repeats, start, frame_num, inter, total, mult=(0, 0)):
    """planner + device generator forms on the CPU for one PLFG programme: (rc, message, closed, running, table, plfg)"""
    g = p.plfgParams
    regs = [0] * (6 + g.maxChirpOrdinalNum + 2 * g.maxNumOfDifferentChirps)
    regs[0], regs[2], regs[3], regs[4], regs[5] = 1, frame_num, inter, len(order), start
    for c, segs in enumerate(chirps):
        regs[6 + c] = len(segs)
    for d, (o, r) in enumerate(zip(order, repeats)):
        regs[6 + g.maxChirpOrdinalNum + d] = r
        regs[6 + g.maxChirpOrdinalNum + g.maxNumOfDifferentChirps + d] = o
    ram = [0] * (g.maxChirpOrdinalNum * g.maxNumOfSegments)
    for c, segs in enumerate(chirps):
        for i, sg in enumerate(segs):
            ram[c * g.maxNumOfSegments + i] = oracle.plfg_ram_word(*sg, output_width=g.outputWidthInt, samples_width=g.maxNumOfSamplesWidth)
    cp = _c_params(p.fftParams, p.magParams, p.cfarParams, p.plfgParams, p.ncoParams)
    closed = np.zeros(total, np.uint32); running = np.zeros(total, np.uint32); table = np.zeros(total, np.uint32); plfg = np.zeros(total, np.int32)
    err = C.create_string_buffer(512)
    rc = emu.rsp_emu_generate(C.byref(cp), _u32(regs).ctypes.data_as(C.c_void_p), len(regs), _u32(ram).ctypes.data_as(C.c_void_p), len(ram),
                              _u32(mult).ctypes.data_as(C.c_void_p), C.c_uint64(total), closed.ctypes.data_as(C.c_void_p),
                              running.ctypes.data_as(C.c_void_p), table.ctypes.data_as(C.c_void_p), plfg.ctypes.data_as(C.c_void_p), err, 512)
    return rc, err.value.decode(), closed, running, table, plfg


def emu_generator_check(emu, oracle, p, chirps, order, repeats, start, frame_num, inter, want_plfg, mult=(0, 0), what=None):
    total = len(want_plfg)
    rc, msg, closed, running, table, plfg = emu_generate(emu, p, oracle, chirps, order, repeats, start, frame_num, inter, total, mult)
    assert rc == 0, (msg, what)
    assert (plfg == want_plfg).all(), what
    n = p.ncoParams
    want, _ = oracle.nco_words(want_plfg, n.tableSize, n.tableWidth, n.phaseWidth, n.phaseAccEnable, use_multiplier=n.useMultiplier,
                               enable_multiplying=bool(mult[0]), multiplying_factor=mult[1], rounding_mode=n.roundingMode, rasterized=n.rasterizedMode)
    assert (closed == want).all() and (running == want).all(), what
    if n.phaseWidth <= 12 and not n.useMultiplier:
        assert (table == want).all(), what      # the fused kernel's beats
    if frame_num:                               # the schedule ends after frameNum frames: one sample more is refused
        assert emu_generate(emu, p, oracle, chirps, order, repeats, start, frame_num, inter, total + 1, mult)[0] == _abi.RSP_EINVAL, what


@pytest.mark.parametrize("nco_kw,plfg_kw", [(dict(), dict()), (dict(phaseAccEnable=False), dict()), (dict(phaseWidth=14, tableSize=1024), dict()),
                                            (dict(phaseWidth=12, tableSize=256, tableWidth=12), dict(outputWidthInt=12, outputWidthFrac=4)),
                                            (dict(useMultiplier=True, roundingMode="Convergent"), dict()),
                                            (dict(rasterizedMode=True, tableSize=100), dict()),                      # phaseWidth 9: beat table
                                            (dict(rasterizedMode=True, tableSize=3000), dict())])                    # phaseWidth 14: look-up in the kernel
@pytest.mark.parametrize("frame_num,inter", [(3, 9), (0, 5), (2, 1)])
def test_emu_generator(emu, oracle, nco_kw, plfg_kw, frame_num, inter):
    """the planner's segment list + the device's closed form, running form (runs of 8 as generator_kernel) and the fused kernel's
    per-period phase table, sample by sample against the oracle PLFG / NCO: four chirp types, repeats, frames, pauses"""
    p = R.rsp_chain_vanilla_app_params()
    for k, v in nco_kw.items():
        setattr(p.ncoParams, k, v)
    for k, v in plfg_kw.items():
        setattr(p.plfgParams, k, v)
    chirps = [[(3, 200, True, False), (5, 120, False, False)], [(1, 255, True, False), (0, 40, False, True), (2, 180, False, False)],
              [(7, 1, True, False)], [(11, 90, False, False), (11, 90, True, False), (4, 33, True, False)]]
    order, repeats, start = [2, 0, 1, 3, 0], [2, 1, 3, 1, 2], 37
    frame_len = 2 * 2 + 320 + 3 * 475 + 213 + 2 * 320
    total = frame_num * frame_len + (frame_num - 1) * max(inter - 1, 0) if frame_num else 3 * frame_len + 100
    want_plfg = oracle_plfg(oracle, p, start, chirps, order, repeats, frame_num, inter).run(total)
    assert len(want_plfg) == total
    emu_generator_check(emu, oracle, p, chirps, order, repeats, start, frame_num, inter, want_plfg, mult=(1, 0x3000) if nco_kw.get("useMultiplier") else (0, 0))


def test_emu_generator_randomised_programmes(emu, oracle):
    """200 seeded random PLFG programmes (tests/util.py:random_plfg_programme) through the planner and the three device forms of the generator on the
    CPU, against the oracle PLFG / NCO; the same programmes run on the GPU in tests/test_gpu_round2.py"""
    from tests.util import random_plfg_programme
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "4242")))
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "200"))):
        nco_kw, chirps, order, repeats, frame_num, inter, start, n_endless = random_plfg_programme(rng)
        p = R.rsp_chain_vanilla_app_params()
        for k, v in nco_kw.items():
            setattr(p.ncoParams, k, v)
        what = (trial, nco_kw, chirps, order, repeats, frame_num, inter, start)
        want_plfg = oracle_plfg(oracle, p, start, chirps, order, repeats, frame_num, inter).run(n_endless if not frame_num else 1 << 18)
        assert 0 < len(want_plfg) < 1 << 18, what
        emu_generator_check(emu, oracle, p, chirps, order, repeats, start, frame_num, inter, want_plfg, what=what)


def test_emu_randomised_run_time_parameter_chain(emu, oracle):
    """150 seeded random configurations of the run-time-parameter kernel's whole parameter space on the CPU (tests/util.py:random_generic_configuration:
    sizes 8 ... 512 and run-time sizes, binPoint 0 ... 14, all trim types, 3 / 4 multipliers, radix 2 / 2^2, DIF / DIT, keep-LSB stages, IFFT through
    the register, every magnitude mode, the four CFAR cores with CASH, sendCut, peak grouping, linear / log): planner + device code bit-exact against
    the oracle, or refused; the same generator runs on the GPU (tests/test_gpu_round2.py)"""
    from tests.util import random_generic_configuration
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "777")))
    ran = refused = 0
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "150"))):
        c = random_generic_configuration(rng, trial, fft_dir_reg=True)
        if c is None:
            continue
        p, log_n, sel, regs, beats, fft_dir, what = c
        try:
            got, _, _, _ = emu_fixed(emu, p, log_n, sel, regs, beats, fft_dir=fft_dir)
        except RuntimeError as e:
            assert "rc=1:" in str(e) or "rc=2:" in str(e), (what, str(e))      # RSP_EINVAL / RSP_EUNSUPPORTED with a message
            refused += 1
            continue
        p.fftParams.fftDir = bool(fft_dir)      # (the oracle takes the direction the register selected)
        fft, mag, cfar = oracle_cfgs(oracle, p, log_n, sel, regs)
        want = oracle.chain_fixed_words(fft, mag, cfar, beats)
        assert (got == want).all(), (what, int((got != want).sum()))
        ran += 1
    assert ran >= 60 and refused <= ran // 10, (ran, refused)


def test_emu_randomised_fp32_cores(emu, oracle):
    """40 seeded random configurations of the DspReal flavour on the CPU (tests/util.py:random_fp32_configuration: the four CFAR cores -- Mem, Lis / GOS,
    ASR / CASH, combined-window OS -- window, guard, modes, ordered-statistic indices, linear / log, peak grouping, 64 / 256 / 1024 points): thresholds
    within 1e-4 of the fp64 dataflow, decided detections identical; the same generator runs on the GPU (tests/test_gpu_round2.py)"""
    from tests.util import check_fp32_against_f64, random_fp32_configuration
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "888")))
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "40"))):
        c = random_fp32_configuration(rng, trial, sizes=(6, 8, 10))
        if c is None:
            continue
        p, log_n, regs, scaler, x32, what = c
        n, rows = 1 << log_n, x32.shape[0]
        cp = _c_params(p.fftParams, p.magParams, p.cfarParams)
        thr = np.zeros((rows, n), np.float32); mask = np.zeros((rows, n // 32), np.uint32)
        err = C.create_string_buffer(512)
        rc = emu.rsp_emu_run_f32(C.byref(cp), _u32([log_n, 1, 0, 0]).ctypes.data_as(C.c_void_p), _u32([2]).ctypes.data_as(C.c_void_p),
                                 cfar_reg_words(p, regs).ctypes.data_as(C.c_void_p), C.c_float(scaler), STAGE_FFT, STAGE_CFAR, x32.ctypes.data_as(C.c_void_p),
                                 C.c_uint32(rows), thr.ctypes.data_as(C.c_void_p), mask.ctypes.data_as(C.c_void_p), err, 512)
        assert rc == 0, (err.value.decode(), what)
        check_fp32_against_f64(oracle, what, log_n, regs, scaler, x32, thr, unpack_mask(mask, n))


@pytest.mark.parametrize("bp,log_n,mode_want", [(0, 10, 4), (1, 8, 4), (2, 6, 5), (12, 10, 5), (14, 8, 5), (15, 5, 5)])
def test_emu_convergent_modes(emu, oracle, bp, log_n, mode_want):
    """trimType = Convergent (the reference's SQNR specs: Radix22SQNRSpec) on the compile-time modes FASTC / FAST3C of the run-time-parameter kernel:
    full-range random beats (every tie and wrap), bit-exact against the oracle, chain and FFT block alone"""
    n = 1 << log_n
    p = chain_params(n, bp, "Convergent", W=max(2, min(16, n // 8)), G=1)
    W = p.cfarParams.leadLaggWindowSize
    regs = dict(fftWin=n, thresholdScaler=int(3.0 * (1 << bp)), logOrLinearMode=1, divSum=max(0, W.bit_length() - 1), peakGrouping=0, cfarMode=0, windowCells=W, guardCells=1)
    beats = np.concatenate([synth.random_beats(6, n, seed=40 + log_n), synth.chirps_fixed(3, n, seed=41 + log_n), np.full((1, n), 0x80008000, dtype=np.uint32)])
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    got, _, kmode, _ = emu_fixed(emu, p, log_n, 2, regs, beats)
    assert kmode == mode_want, kmode
    assert (got == oracle.chain_fixed_words(fft, mag, cfar, beats)).all()
    got, _, kmode, _ = emu_fixed(emu, p, log_n, 2, regs, beats, last=STAGE_FFT)
    assert kmode == mode_want and (got == oracle.fft_fixed_words(fft, beats)).all()
