"""GPU parity, part 3 (through the C ABI): configurations that were built but untested after round 1 (VERDICT r01 "what's weak" 2-3):
sendCut, LogMagMux bypass through the chain, the detection-list-only mode of the fast kernel, the fp32 flavour with the Lis / ASR cores,
peak grouping and the other magnitude modes, C3 at its full row count, and the generator beyond one chirp type."""
import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import synth
from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, GOSCFARType,
                                    MAGParams)
from oracle.glue import f64_chain, oracle_plfg, unpack_mask, write_plfg
from tests.util import configure, decode_cfar_words, from_dev_u32, oracle_cfgs, to_dev

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4


def fixed_params(n, W, G, bp=0, ctype=CACFARType, send_cut=False, cash=False):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=bp),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=bp, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, bp), protoThreshold=FixedPoint(16, bp), protoScaler=FixedPoint(16, 0),
                              leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype, sendCut=send_cut,
                              includeCASH=cash, minSubWindowSize=4 if cash else None))


def run_fixed(cuda, p, log_n, sel, regs, beats, **kw):
    n = 1 << log_n
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, sel, regs)
        d_out = cuda.empty((beats.shape[0], n), dtype=cuda.int32, device="cuda")
        ch.run(to_dev(cuda, beats), beats.shape[0], 1, out=d_out, **kw)
        cuda.cuda.synchronize()
        return from_dev_u32(d_out), ch.variant


@pytest.mark.parametrize("ctype,cash", [(CACFARType, False), (GOSCFARType, False), (CACFARType, True)])
@pytest.mark.parametrize("log_n", [9, 10])
def test_send_cut_outputs_the_cell_under_test(cuda, oracle, log_n, ctype, cash):
    """CFARParams.sendCut = true (CFARDspBlock.scala:78-79): the output beat is the cell under test, whatever the core"""
    n = 1 << log_n
    p = fixed_params(n, 16, 2, ctype=ctype, send_cut=True, cash=cash)
    regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=2, indexLagg=12, indexLead=5,
                subWindowSize=8)
    beats = np.concatenate([synth.chirps_fixed(9, n, seed=log_n), synth.random_beats(3, n, seed=3)])
    got, variant = run_fixed(cuda, p, log_n, 2, regs, beats)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, _, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    assert (got == want).all(), variant
    assert (got == (want_mag & 0xFFFF)).all()        # the CUT is the magnitude beat's low half


@pytest.mark.parametrize("log_n", [6, 10, 12])
def test_mag_bypass_through_the_chain(cuda, oracle, log_n):
    """LogMagMux sel = 0 (bypass, LogMagMux.scala:68-90) inside the chain: the CFAR sees the low 16 bits (imaginary part) of the FFT beat"""
    n = 1 << log_n
    W = 8 if log_n == 6 else 16
    p = fixed_params(n, W, 2)
    regs = dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=1, windowCells=W, guardCells=2)
    beats = np.concatenate([synth.chirps_fixed(6, n, seed=log_n + 40), synth.random_beats(4, n, seed=log_n)])
    got, variant = run_fixed(cuda, p, log_n, 0, regs, beats)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 0, regs)
    want, want_fft, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    assert (want_mag == want_fft).all()
    assert (got == want).all(), variant


@pytest.mark.parametrize("log_n", [10, 14])
def test_fast_kernel_detection_list_only(cuda, oracle, log_n):
    """rsp_run with d_out = NULL: the fast kernel builds no output beats and reports the same detections as the full run"""
    n = 1 << log_n
    W, G = (16, 4) if log_n == 10 else (64, 8)
    p = fixed_params(n, W, G)
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    rows = 77 if log_n == 10 else 7
    beats = synth.chirps_fixed(rows, n, seed=5 + log_n, noise=2.0 ** -7 if log_n == 10 else 2.0 ** -2)   # a noise floor above 1 LSB after the 1/N
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, _, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    thr, bins, peak = decode_cfar_words(want, n)
    assert peak.sum() < 1 << 16
    exp = {(r, b): (int(thr[r, b]), int(np.int16(want_mag[r, b] & 0xFFFF))) for r, b in np.argwhere(peak == 1)}
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>"
        det_list = cuda.zeros((1 << 16, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        l0 = ch.kernel_launches
        ch.run(to_dev(cuda, beats), rows, 1, out=None, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        assert ch.kernel_launches == l0 + 1
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
        # the host entry point with h_out = NULL
        n_found = ch.run_host_dets_raw(np.ascontiguousarray(beats).ctypes.data, rows, 1, 1 << 16)
        host = np.frombuffer(ch.last_detections, dtype=np.int32).reshape(-1, 4)[:n_found].copy()
    assert k == len(exp) and k > 0
    assert {(int(d[0]), int(d[1])): (int(d[2]) & 0xFFFF, int(d[3])) for d in dets} == exp
    assert n_found == k and {(int(d[0]), int(d[1])): (int(d[2]) & 0xFFFF, int(d[3])) for d in host} == exp


def test_peak_grouping_without_guard_cells_is_refused_on_lis_and_asr(cuda):
    """the reference substitutes the window ends for the missing guard cells (CFARCoreWithLis.scala:191-192); not built -> RSP_EUNSUPPORTED"""
    p = fixed_params(1024, 16, 2, ctype=GOSCFARType)
    regs = dict(fftWin=1024, thresholdScaler=3, logOrLinearMode=1, peakGrouping=1, cfarMode=0, windowCells=16, guardCells=0, indexLagg=8, indexLead=8)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        d = cuda.zeros((2, 1024), dtype=cuda.int32, device="cuda")
        with pytest.raises(R.RspError) as e:
            ch.run(d, 2, 1, out=d.clone())
        assert e.value.status == R._abi.RSP_EUNSUPPORTED


# ---------------------------------------------------------------------------------------------------- fp32 flavour
def f32_params(n, W, G, ctype=CACFARType, cash=False):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True), magParams=MAGParams.fixed(),
        cfarParams=CFARParams(leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype, sendCut=False,
                              includeCASH=cash, minSubWindowSize=4 if cash else None))


def run_f32(cuda, p, log_n, sel, regs, scaler, x32):
    n, rows = 1 << log_n, x32.shape[0]
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, sel, regs)
        ch.set_threshold_scaler_f32(scaler)
        d_thr = cuda.empty((rows, n), dtype=cuda.float32, device="cuda")
        d_mask = cuda.zeros((rows, max(1, n // 32)), dtype=cuda.int32, device="cuda")
        ch.run(cuda.from_numpy(x32).cuda(), rows, 1, out=d_thr, peak_mask=d_mask)
        cuda.cuda.synchronize()
        return d_thr.cpu().numpy(), unpack_mask(d_mask.cpu().numpy().view(np.uint32), n), ch.variant


def check_f32(thr, pk, mag, want_thr, want_pk, what, tol=REL_TOL):
    scale = np.abs(want_thr).max()
    rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale)
    assert rel.max() < tol, f"{what}: max relative threshold error {rel.max():.3e}"
    sure = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale) > 10 * REL_TOL
    assert (pk[sure] == want_pk[sure]).all(), f"{what}: detections differ"
    assert sure.mean() > 0.99


@pytest.mark.parametrize("core", ["gos", "gos-ca", "cash", "asr-ca"])
@pytest.mark.parametrize("pg", [0, 1])
def test_fp32_lis_and_asr_cores(cuda, oracle, core, pg):
    """the DspReal flavour with CFARCoreWithLis (ordered statistic and its cell-averaging mode) and CFARCoreWithASR (CASH and the
    sub-window averages), with and without peak grouping: thresholds zero unless both windows are full (CFARCoreWithLis.scala:172-173)"""
    log_n, W, G = 10, 16, 2
    n = 1 << log_n
    xc = synth.chirps_complex(6, n, seed=17)
    xc[2] *= 1e-2
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=n, logOrLinearMode=1, divSum=4, peakGrouping=pg, cfarMode=0, windowCells=W, guardCells=G, indexLagg=12, indexLead=6, subWindowSize=4)
    kw = {}
    if core == "gos":
        p, kw = f32_params(n, W, G, GOSCFARType), dict(core=oracle.CORE_LIS, index_lagg=12, index_lead=6)
    elif core == "gos-ca":
        p, kw = f32_params(n, W, G, R.GOSCACFARType), dict(core=oracle.CORE_LIS)
        regs["cfarAlgorithm"] = 0
    elif core == "cash":
        p, kw = f32_params(n, W, G, cash=True), dict(core=oracle.CORE_ASR, sub_cells=4)
        regs["cfarMode"] = 3
        regs["divSum"] = 2
    else:
        p, kw = f32_params(n, W, G, cash=True), dict(core=oracle.CORE_ASR, sub_cells=4)
        regs["cfarMode"] = 1
        regs["divSum"] = 2
    for sel in (2, 1):
        thr, pk, variant = run_f32(cuda, p, log_n, sel, regs, 3.0, x32)
        assert variant == f"chain_f32<{log_n}>", variant
        mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, sel, regs, 3.0, W, G, **kw)
        if sel == 1:    # power cells: compared in the amplitude domain (sqrt), where the 1e-4 of the fp32 transform is defined -- the error of
            #              an fp32 FFT is absolute (eps * log2 N * the row's largest bin), and squaring doubles the relative error of every cell
            check_f32(np.sqrt(thr), pk, np.sqrt(mag), np.sqrt(want_thr), want_pk, f"{core} pg {pg} sel {sel}")
        else:
            check_f32(thr, pk, mag, want_thr, want_pk, f"{core} pg {pg} sel {sel}")
        assert (want_thr[:, :W + G] == 0).all() and (thr[:, :W + G] == 0).all()


@pytest.mark.parametrize("sel", [1, 3])
@pytest.mark.parametrize("pg", [0, 1])
def test_fp32_mem_core_other_magnitudes_and_peak_grouping(cuda, oracle, sel, pg):
    """|z|^2 and log2 magnitude, peak grouping, on the Mem core (log mode for the log2 magnitude: threshold = mean + scaler)"""
    log_n, W, G = 11, 32, 4
    n = 1 << log_n
    xc = synth.chirps_complex(5, n, seed=23)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=n, logOrLinearMode=0 if sel == 3 else 1, divSum=5, peakGrouping=pg, cfarMode=0, windowCells=W, guardCells=G)
    scaler = 2.0 if sel == 3 else 6.0
    thr, pk, variant = run_f32(cuda, f32_params(n, W, G), log_n, sel, regs, scaler, x32)
    assert variant.startswith("chain_f32")
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, sel, regs, scaler, W, G)
    if sel == 3:      # log2 of a magnitude known to 1e-6 relative: absolute error ~1e-6 on values of order 10
        assert np.abs(thr - want_thr).max() < 1e-4 * np.abs(want_thr).max()
        sure = np.abs(mag - want_thr) > 1e-3
        assert (pk[sure] == want_pk[sure]).all()
    else:
        check_f32(thr, pk, mag, want_thr, want_pk, f"sel {sel} pg {pg}")
    assert want_pk.sum() >= 5


def test_c3_full_frame_fp32(cuda, oracle):
    """BASELINE config 2 at its full size: 4 RX x 512 chirps x 2048 points, fp32, GO-CFAR (W 32, G 4).  The fp64 dataflow checks the 16
    distinct rows; the size-independent property is row independence: every replica of the 16-row block equals the first."""
    n, log_n, W, G, n_rx, n_chirps = 2048, 11, 32, 4, 4, 512
    rows = n_rx * n_chirps
    xc = synth.chirps_complex(16, n, seed=3, rx_phase=np.repeat(np.arange(4) * np.pi / 8, 4))
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=n, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=1, windowCells=W, guardCells=G)
    with R.FftMagCfarChainVanilla(f32_params(n, W, G)) as ch:
        configure(ch, log_n, 2, regs)
        ch.set_threshold_scaler_f32(3.5)
        assert ch.variant == "chain_f32_fast<11>"
        d_in = cuda.from_numpy(x32).cuda().repeat(rows // 16, 1, 1).contiguous()
        d_thr = cuda.empty((rows, n), dtype=cuda.float32, device="cuda")
        d_mask = cuda.zeros((rows, n // 32), dtype=cuda.int32, device="cuda")
        ch.run(d_in, n_chirps, n_rx, out=d_thr, peak_mask=d_mask)
        cuda.cuda.synchronize()
        assert bool((d_thr.view(rows // 16, 16, n) == d_thr[:16].unsqueeze(0)).all().item())
        assert bool((d_mask.view(rows // 16, 16, n // 32) == d_mask[:16].unsqueeze(0)).all().item())
        thr = d_thr[:16].cpu().numpy()
        pk = unpack_mask(d_mask[:16].cpu().numpy().view(np.uint32), n)
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, 3.5, W, G)
    check_f32(thr, pk, mag, want_thr, want_pk, "C3")
    assert (pk == want_pk).all() and want_pk.sum() >= 32


# ---------------------------------------------------------------------------------------------------- PLFG -> NCO generator
CHIRPS = [[(3, 200, True, False), (5, 120, False, False)],                       # chirp 0: up, down
          [(1, 255, True, False), (0, 40, False, True), (2, 180, False, False)],  # chirp 1: up, reset plateau, down
          [(7, 1, True, False)],                                                  # chirp 2: nSamples 1 is bumped to 2
          [(11, 90, False, False), (11, 90, True, False), (4, 33, True, False)]]  # chirp 3: down, up, up


@pytest.mark.parametrize("nco_kw,plfg_kw", [
    (dict(), dict()),                                                   # RspChainVanillaApp: phaseWidth 9, full-wave beat table
    (dict(phaseAccEnable=False), dict()),                               # the PLFG word is the phase
    (dict(phaseWidth=14, tableSize=1024), dict()),                      # phaseWidth > 12: the in-kernel quarter-wave look-up (nco_lookup)
    (dict(phaseWidth=14, tableSize=256, tableWidth=12), dict()),        # phase bits below the table address are dropped
    (dict(phaseWidth=14, tableSize=1024), dict(outputWidthInt=12, outputWidthFrac=4)),   # config word sliced at outputWidthInt
])
@pytest.mark.parametrize("frame_num,inter", [(3, 9), (0, 5), (2, 1)])
def test_generator_many_chirp_types_frames_and_pauses(cuda, oracle, nco_kw, plfg_kw, frame_num, inter):
    """four chirp types addressed through chirpOrdinalNums, repeated chirps, finite frameNum crossing two inter-frame pauses (RTL
    pause length, PLFGDspBlockMem.scala:242-275) and the period wrap of an endless schedule; PLFG words and NCO beats bit-exact, then
    the chain fed by the generator (rsp_run_from_generator)"""
    p = R.rsp_chain_vanilla_app_params()
    for k, v in nco_kw.items():
        setattr(p.ncoParams, k, v)
    for k, v in plfg_kw.items():
        setattr(p.plfgParams, k, v)
    order, repeats = [2, 0, 1, 3, 0], [2, 1, 3, 1, 2]
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    start = 37
    op = oracle_plfg(oracle, p, start, CHIRPS, order, repeats, frame_num, inter)
    frame_len = 2 * 2 + 320 + 3 * 475 + 213 + 2 * 320
    total = frame_num * frame_len + (frame_num - 1) * max(inter - 1, 0) if frame_num else 6 * frame_len + 100
    n_rows = min(total // 1024, 6)
    with R.RspChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        write_plfg(ch, p, oracle, start, CHIRPS, order, repeats, frame_num, inter)
        d_nco = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        d_plfg = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        ch.generate(total, d_nco, d_plfg)
        d_out = cuda.zeros((n_rows, 1024), dtype=cuda.int32, device="cuda")
        ch.run_from_generator(n_rows, 1, out=d_out)
        cuda.cuda.synchronize()
        got_nco, got_plfg, got_out = from_dev_u32(d_nco), d_plfg.cpu().numpy(), from_dev_u32(d_out)
        if frame_num:   # the schedule ends after frameNum frames: one sample more is refused
            with pytest.raises(R.RspError):
                ch.generate(total + 1, cuda.zeros(total + 1, dtype=cuda.int32, device="cuda"), None)
    want_plfg = op.run(total)
    assert len(want_plfg) == total
    assert (got_plfg == want_plfg).all(), f"first difference at {np.argwhere(got_plfg != want_plfg)[:3].ravel()}"
    n = p.ncoParams
    want_nco, _ = oracle.nco_words(want_plfg, n.tableSize, n.tableWidth, n.phaseWidth, n.phaseAccEnable)
    assert (got_nco == want_nco).all(), f"first difference at {np.argwhere(got_nco != want_nco)[:3].ravel()}"
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want_out = oracle.chain_fixed_words(fft, mag, cfar, want_nco[:n_rows * 1024].reshape(n_rows, 1024))
    assert (got_out == want_out).all()


def test_generator_refuses_a_pause_the_hardware_never_leaves(cuda, oracle):
    """framePauseCounter is UInt(6.W): interframeNumOfSamples >= 64 never restarts the frame (PLFGDspBlockMem.scala:144,262)"""
    p = R.rsp_chain_vanilla_app_params()
    with R.RspChainVanilla(p) as ch:
        write_plfg(ch, p, oracle, 0, CHIRPS[:1], frame_num=2, inter=64)
        with pytest.raises(R.RspError) as e:
            ch.generate(100, cuda.zeros(100, dtype=cuda.int32, device="cuda"), None)
        assert e.value.status == R._abi.RSP_EUNSUPPORTED


# ---------------------------------------------------------------------------------------------------- C4 reading (B): golden cfarOS
def test_c4_reading_b_combined_window_os(cuda, oracle):
    """BASELINE configs[3], reading (B) of "ref window 64, k = 48": the Scala golden cfarOS (CFARUtils.scala:159-196) with
    referenceCells = 32 per side -> 64 cells in one sorted window, ascending index floor(1.5 * 32) = 48.  Full chain on 4096-point rows
    (bit-exact against the oracle's integer core, itself equal to the golden, tests/test_oracle_cfar.py) and CFAR alone on stored
    16-bit magnitude profiles (exponential clutter + targets) checked against the golden directly."""
    n, log_n, Rc, G = 4096, 12, 32, 4
    p = fixed_params(n, 64, G, ctype=R.OSGoldenCFARType)
    regs = dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=Rc, guardCells=G)
    beats = np.concatenate([synth.chirps_fixed(5, n, seed=44), synth.random_beats(2, n, seed=45)])
    got, variant = run_fixed(cuda, p, log_n, 2, regs, beats)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    assert (got == want).all(), variant
    # CFAR alone on stored profiles
    rng = np.random.default_rng(4)
    prof = np.floor(rng.exponential(256.0, (6, n))).astype(np.int32)
    prof[:, [300, 1200, 2500, 3900]] += 4096
    pw = fixed_params(n, 64, G, ctype=R.OSGoldenCFARType)
    pw.cfarParams.protoThreshold = FixedPoint(18, 0)       # room for 2 * (clutter + target)
    with R.FftMagCfarChainVanilla(pw) as ch:
        configure(ch, log_n, 2, regs)
        d_out = cuda.empty((6, n), dtype=cuda.int32, device="cuda")
        ch.cfar(to_dev(cuda, prof.astype(np.uint32)), 6, out=d_out)
        cuda.cuda.synchronize()
        out = from_dev_u32(d_out)
    thr, bins, peak = out >> 13, (out >> 1) & 4095, out & 1
    for r in range(6):
        g_thr, g_peaks = oracle.golden_cfar_os(prof[r].astype(np.float64), Rc, G, 2.0)
        assert (thr[r] == g_thr.astype(np.int64)).all()
        assert np.flatnonzero(peak[r]).tolist() == sorted(set(g_peaks.tolist()))
    assert (bins == np.arange(n)[None, :]).all() and peak.sum() >= 24


def test_fp32_combined_window_os(cuda, oracle):
    """the same statistic in the DspReal flavour: the threshold is one of the magnitudes times the scaler"""
    log_n, W, G = 11, 32, 4
    n = 1 << log_n
    xc = synth.chirps_complex(4, n, seed=29)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=n, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    thr, pk, variant = run_f32(cuda, f32_params(n, 64, G, R.OSGoldenCFARType), log_n, 2, regs, 2.5, x32)
    assert variant == f"chain_f32<{log_n}>"
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, 2.5, W, G, core=oracle.CORE_OS)
    check_f32(thr, pk, mag, want_thr, want_pk, "fp32 OS")
    for r in range(4):
        g_thr, g_peaks = oracle.golden_cfar_os(mag[r], W, G, 2.5)
        assert np.allclose(want_thr[r], g_thr, rtol=0, atol=0)


# ---------------------------------------------------------------------------------------------------- wide FFT block, grow stages, overflow status
def pack_beats64(re, im, w):
    m = np.uint64((1 << w) - 1)
    return ((np.asarray(re, np.int64).astype(np.uint64) & m) << np.uint64(w)) | (np.asarray(im, np.int64).astype(np.uint64) & m)


def wide_params(n, dw, bp, tw, grow=None, keep=None, **kw):
    s = n.bit_length() - 1
    fp = FFTParams.fixed(dataWidth=dw, twiddleWidth=tw, numPoints=n, binPoint=bp, expandLogic=list(grow or [0] * s),
                         keepMSBorLSB=list(keep or [True] * s), **kw)
    return FftMagCfarVanillaParameters(fftParams=fp, magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
                                       cfarParams=CFARParams(protoIn=FixedPoint(16, 0), protoThreshold=FixedPoint(16, 0), protoScaler=FixedPoint(16, 0),
                                                             leadLaggWindowSize=16, guardWindowSize=2, fftSize=max(n, 64), CFARAlgorithm=CACFARType))


@pytest.mark.parametrize("name,n,dw,bp,tw,grow,kw", [
    ("24-bit scaled", 1024, 24, 22, 24, None, dict(trimType="Convergent", useBitReverse=True)),
    ("16-bit grow", 1024, 16, 14, 16, [1] * 10, dict(trimType="Convergent", useBitReverse=True)),
    ("24-bit grow (128-bit beats)", 1024, 24, 22, 24, [1] * 10, dict(trimType="Convergent", useBitReverse=True)),
    ("20-bit, some stages grow, one keeps LSBs, 4 multipliers, bit-reversed out", 256, 20, 10, 18, [1, 0, 0, 1, 0, 1, 0, 0],
     dict(trimType="RoundHalfUp", use4Muls=True, useBitReverse=False, keepMSBorLSB=[True, True, False, True, True, True, True, True])),
    ("32-bit data, floor", 64, 32, 20, 32, None, dict(trimType="RoundDown", useBitReverse=True)),
    ("24-bit run-time size 256 of 4096, grow in the skipped stages", 4096, 24, 12, 20, [1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 0, 0],
     dict(trimType="RoundHalfUp", useBitReverse=True, runTime=True)),
    ("8192-point 17-bit", 8192, 17, 8, 16, None, dict(trimType="Convergent", useBitReverse=True)),
])
def test_wide_fft_block(cuda, oracle, name, n, dw, bp, tw, grow, kw):
    """AXI4FFTBlock with beats wider than Cat(re16, im16): the configurations behind the reference's 24-bit and grow SQNR results
    (Radix22SQNRSpec.scala:60-150).  Bit-exact against the oracle (which reproduces the published SQNR points,
    tests/test_oracle_fft_mag.py), full-range random inputs, overflow status included."""
    keep = kw.pop("keepMSBorLSB", None)
    p = wide_params(n, dw, bp, tw, grow, keep, overflowReg=True, **kw)
    log_n = 8 if "run-time" in name else n.bit_length() - 1
    m = 1 << log_n
    rng = np.random.default_rng(len(name))
    rows = 5
    re = rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), (rows, m)); im = rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), (rows, m))
    re[1] >>= 3; im[1] >>= 3                    # a row that stays clear of the wrap corners
    beats = pack_beats64(re, im, dw)
    fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, dict(fftWin=max(m, 64)))
    want, want_ovf = oracle.fft_fixed_words64(fft, beats)
    with R.FftMagCfarChainVanilla(p) as ch:
        if p.fftParams.runTime:
            ch.memWriteWord(p.fftAddress.base, log_n)
        assert ch.fft_in_beat_bytes == 8
        k = ch.fft_out_beat_bytes // 8
        assert k == (2 if want.ndim == 3 else 1)
        d_in = cuda.from_numpy(beats.view(np.int64)).cuda()
        d_out = cuda.zeros((rows, m, k), dtype=cuda.int64, device="cuda")
        ch.fft(d_in, rows, d_out)
        got = d_out.cpu().numpy().view(np.uint64).reshape(want.shape)
        ovf = ch.fft_overflow_flags()
        with pytest.raises(R.RspError):          # LogMagMux / CFAR take 16-bit cells: the chain refuses wide beats
            ch.run(d_in, rows, 1, out=d_out)
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"{name}: {bad.shape[0]} beats differ, first at {bad[0]}"
    smax = n.bit_length() - 1
    assert ovf == want_ovf << (smax - log_n), f"{name}: overflow status {ovf:#x} vs {want_ovf << (smax - log_n):#x}"


def test_wide_fft_window_and_ifft(cuda, oracle):
    """window multiply in front of the wide transform (in * window truncated on assignment, SDFFFT.scala:233-234) and the IFFT of a
    24-bit chain without grow (output scalar 1.0); IFFT with grow stages is refused (the 1/N literal of SDFChainRadix22.scala:362)"""
    from rsp_chains_b200 import WindowFunctionTypes as WT
    n, dw, bp, tw = 512, 24, 20, 24
    rng = np.random.default_rng(5)
    re = rng.integers(-(1 << 22), 1 << 22, (3, n)); im = rng.integers(-(1 << 22), 1 << 22, (3, n))
    for fwd in (True, False):
        p = wide_params(n, dw, bp, tw, trimType="RoundHalfUp", useBitReverse=True, fftDir=fwd, windowFunc=WT.Hanning(16))
        fft, _, _ = oracle_cfgs(oracle, p, 9, 2, dict(fftWin=n))
        wre, wim = re, im
        if fwd:        # the window acts on the forward transform only (SDFFFT.scala:229-245)
            raw = oracle.window_raw(oracle.window_coeffs(2, n), 16).astype(np.int64)
            wrap = lambda v: ((v + (1 << (dw - 1))) & ((1 << dw) - 1)) - (1 << (dw - 1))   # noqa: E731
            wre, wim = wrap((re * raw[None, :]) >> 14), wrap((im * raw[None, :]) >> 14)
        want, _ = oracle.fft_fixed_words64(fft, pack_beats64(wre, wim, dw))
        with R.FftMagCfarChainVanilla(p) as ch:
            d_out = cuda.zeros((3, n), dtype=cuda.int64, device="cuda")
            ch.fft(cuda.from_numpy(pack_beats64(re, im, dw).view(np.int64)).cuda(), 3, d_out)
            got = d_out.cpu().numpy().view(np.uint64)
        assert (got == want).all(), f"fftDir {fwd}"
    p = wide_params(n, dw, bp, tw, [1] * 9, fftDir=False)
    with R.FftMagCfarChainVanilla(p) as ch:
        with pytest.raises(R.RspError) as e:
            ch.fft(cuda.zeros((1, n), dtype=cuda.int64, device="cuda"), 1, cuda.zeros((1, n, 2), dtype=cuda.int64, device="cuda"))
        assert e.value.status == R._abi.RSP_EUNSUPPORTED


@pytest.mark.parametrize("radix,dit", [("2^2", False), ("2", False), ("2", True), ("2^2", True)])
def test_fft_overflow_status_register(cuda, oracle, radix, dit):
    """FFTParams.overflowReg (FFTBlock.scala:67,81-82,97; SDFChainRadix22.scala:473-486): the per-stage status of the last call through
    rsp_fft_overflow_flags (all stages) and through the register's log2Ceil(numStages)-bit field; rows below half scale report nothing,
    a run-time size reports on the physical stages it uses; the chain (rsp_run) reports it too"""
    n, smax = 1024, 10
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    p.fftParams.binPoint = 0
    p.magParams.binPoint = 0
    p.cfarParams.protoIn = p.cfarParams.protoThreshold = p.cfarParams.protoScaler = FixedPoint(16, 0)
    p.fftParams.overflowReg = True
    p.fftParams.sdfRadix = radix
    p.fftParams.decimType = R.DITDecimType if dit else R.DIFDecimType
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    quiet = synth.chirps_fixed(4, n, seed=9)
    loud = synth.random_beats(4, n, seed=10)
    dc = np.full((2, n), (20000 << 16) & 0xFFFFFFFF, dtype=np.uint32)          # re = 20000: the DC path flags every stage
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        for beats in (quiet, loud, dc):
            fft, _, _ = oracle_cfgs(oracle, p, 10, 2, regs)
            want_out, want_ovf = oracle.fft_fixed_words_ovf(fft, beats)
            d_out = cuda.empty((beats.shape[0], n), dtype=cuda.int32, device="cuda")
            ch.fft(to_dev(cuda, beats), beats.shape[0], d_out)
            assert (from_dev_u32(d_out) == want_out).all()
            assert ch.fft_overflow_flags() == want_ovf
            assert ch.memReadWord(p.fftAddress.base + 4 * p.beatBytes) == want_ovf & 0xF      # log2Ceil(10) = 4 bits
            ch.run(to_dev(cuda, beats), beats.shape[0], 1, out=d_out)
            assert ch.fft_overflow_flags() == want_ovf
        assert oracle.fft_fixed_words_ovf(fft, quiet)[1] == 0 and oracle.fft_fixed_words_ovf(fft, dc)[1] != 0
        if not dit and radix == "2^2":      # run-time size 256: the last 8 physical stages
            configure(ch, 8, 2, dict(regs, fftWin=256))
            fft, _, _ = oracle_cfgs(oracle, p, 8, 2, dict(regs, fftWin=256))
            want_out, want_ovf = oracle.fft_fixed_words_ovf(fft, dc[:, :256])
            d_out = cuda.empty((2, 256), dtype=cuda.int32, device="cuda")
            ch.fft(to_dev(cuda, np.ascontiguousarray(dc[:, :256])), 2, d_out)
            assert ch.fft_overflow_flags() == want_ovf << 2 and want_ovf != 0


# ---------------------------------------------------------------------------------------------------- FFTBlockWithWindowing
@pytest.mark.parametrize("n,bp,kw", [(1024, 0, dict()), (1024, 12, dict(trimType="Convergent")), (256, 0, dict(sdfRadix="2")),
                                      (2048, 0, dict(fftDir=False)), (512, 18, dict(dataWidth=24, twiddleWidth=20))])
def test_fft_block_with_windowing(cuda, oracle, n, bp, kw):
    """FFTBlockWithWindowing (FFTBlockWithWindowing.scala:18-176) driven as the reference's tester drives it
    (FFTBlockWithWindowingSpec.scala:100-110): doWindowing = 1 at csr word 2, then the window as raw protoWin words
    `c * 2^(w - 2)` into the RAM, then the stream.  Bit-exact against the oracle on windowed beats (in both directions: the RAM
    multiplier sits in front of the core and is not gated by fftDir, :159-166), the reference's |delta| <= 4 acceptance against
    fft(x * w) / N (TesterUtil.scala:222-228), a smaller run-time size (the first n coefficients), and doWindowing = 0."""
    from rsp_chains_b200.params import AddressSet
    log_n = n.bit_length() - 1
    dw = kw.get("dataWidth", 16)
    fp = FFTParams.fixed(**{**dict(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=bp), **kw})
    csr, ram = AddressSet(0x010000, 0xFF), AddressSet(0x000000, 0x3FFF)
    coeffs = oracle.window_coeffs(2, n)                     # Hanning
    raw = [int(c * (1 << 14)) for c in coeffs]              # BigDecimal(c * (1 << (w - 2))).toBigInt: truncation (:98)
    amp = 1 << (dw - 3)
    rng = np.random.default_rng(n)
    t = np.arange(n)
    tone = np.round(amp * np.exp(2j * np.pi * (n // 8) * t / n))
    re = np.stack([tone.real.astype(np.int64), rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), n), rng.integers(-amp, amp, n)])
    im = np.stack([tone.imag.astype(np.int64), rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), n), rng.integers(-amp, amp, n)])
    wide = dw > 16
    wrap = lambda v: ((v + (1 << (dw - 1))) & ((1 << dw) - 1)) - (1 << (dw - 1))   # noqa: E731

    def beats_of(r, i, m):
        if wide:
            return pack_beats64(r[:, :m], i[:, :m], dw)
        return (((r[:, :m] & 0xFFFF) << 16) | (i[:, :m] & 0xFFFF)).astype(np.uint32)

    with R.FFTBlockWithWindowing(csr, ram, fp, beatBytes=4) as blk:
        blk.memWriteWord(csr.base + 2 * 4, 1)                # enable windowing
        for i, v in enumerate(raw):
            blk.memWriteWord(ram.base + i * 4, v)
        for m_log in (log_n, log_n - 2):
            m = 1 << m_log
            blk.memWriteWord(csr.base, m_log)
            if kw.get("sdfRadix") != "2" and (log_n - m_log) & 1:
                continue
            w = np.array(raw[:m], dtype=np.int64)
            wre, wim = wrap((re[:, :m] * w) >> 14), wrap((im[:, :m] * w) >> 14)
            pp = FftMagCfarVanillaParameters(fftParams=fp, magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
                                             cfarParams=CFARParams(leadLaggWindowSize=16, guardWindowSize=2, fftSize=n))
            fft, _, _ = oracle_cfgs(oracle, pp, m_log, 2, dict(fftWin=m))
            if wide:
                want, _ = oracle.fft_fixed_words64(fft, pack_beats64(wre, wim, dw))
                d_out = cuda.zeros((3, m), dtype=cuda.int64, device="cuda")
                blk.fft(cuda.from_numpy(beats_of(re, im, m).view(np.int64)).cuda(), 3, d_out)
                got = d_out.cpu().numpy().view(np.uint64)
            else:
                want = oracle.fft_fixed_words(fft, beats_of(wre, wim, m))
                d_out = cuda.zeros((3, m), dtype=cuda.int32, device="cuda")
                blk.fft(to_dev(cuda, beats_of(re, im, m)), 3, d_out)
                got = from_dev_u32(d_out)
            assert (got == want).all(), f"size {m}"
            if not wide and fp.fftDir and m == n and bp == 0:      # the reference's acceptance on the tone (checkFFTError: |delta| <= 4)
                ore, oim = oracle.unpack_iq(got[0])
                ref = np.fft.fft((re[0] + 1j * im[0]) * np.array(raw) / (1 << 14)) / n
                assert np.abs(ore - ref.real).max() <= 4 and np.abs(oim - ref.imag).max() <= 4
        # doWindowing = 0: the plain transform
        blk.memWriteWord(csr.base, log_n)
        blk.memWriteWord(csr.base + 2 * 4, 0)
        assert blk.memReadWord(csr.base + 2 * 4) == 0
        pp = FftMagCfarVanillaParameters(fftParams=fp, magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
                                         cfarParams=CFARParams(leadLaggWindowSize=16, guardWindowSize=2, fftSize=n))
        fft, _, _ = oracle_cfgs(oracle, pp, log_n, 2, dict(fftWin=n))
        if wide:
            want, _ = oracle.fft_fixed_words64(fft, pack_beats64(re, im, dw))
            d_out = cuda.zeros((3, n), dtype=cuda.int64, device="cuda")
            blk.fft(cuda.from_numpy(beats_of(re, im, n).view(np.int64)).cuda(), 3, d_out)
            assert (d_out.cpu().numpy().view(np.uint64) == want).all()
        else:
            want = oracle.fft_fixed_words(fft, beats_of(re, im, n))
            d_out = cuda.zeros((3, n), dtype=cuda.int32, device="cuda")
            blk.fft(to_dev(cuda, beats_of(re, im, n)), 3, d_out)
            assert (from_dev_u32(d_out) == want).all()


def test_fft_block_with_windowing_requires_window_none(cuda):
    from rsp_chains_b200 import WindowFunctionTypes as WT
    from rsp_chains_b200.params import AddressSet
    with pytest.raises(AssertionError):
        R.FFTBlockWithWindowing(AddressSet(0x010000, 0xFF), AddressSet(0, 0xFFF), FFTParams.fixed(numPoints=64, windowFunc=WT.Hanning()))


@pytest.mark.parametrize("rounding", ["RoundHalfUp", "Convergent", "RoundDown"])
def test_nco_output_multiplier(cuda, oracle, rounding):
    """FixedNCOParams.useMultiplier: the two NCO registers (enableMultiplying, multiplyingFactor; NCOLazyModuleBlock.scala:262-269) and
    the product of :354-364, bit-exact against the oracle, disabled / 0.25 (the factor of NCOMultiplierSpec.scala:100) / negative / > 1"""
    p = R.rsp_chain_vanilla_app_params()
    p.ncoParams.useMultiplier = True
    p.ncoParams.roundingMode = rounding
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    n = 3 * 1024
    with R.RspChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        write_plfg(ch, p, oracle, 3, CHIRPS[:2], frame_num=0, inter=7)
        want_plfg = oracle_plfg(oracle, p, 3, CHIRPS[:2], frame_num=0, inter=7).run(n)
        for enable, factor in ((0, 0x1000), (1, 0x1000), (1, 0xA000), (1, 0x7FFF), (1, 0x4000)):
            ch.memWriteWord(p.ncoAddress.base + 1 * p.beatBytes, factor)
            ch.memWriteWord(p.ncoAddress.base, enable)
            assert ch.memReadWord(p.ncoAddress.base + p.beatBytes) == factor
            d_nco = cuda.zeros(n, dtype=cuda.int32, device="cuda")
            ch.generate(n, d_nco, None)
            d_out = cuda.zeros((3, 1024), dtype=cuda.int32, device="cuda")
            ch.run_from_generator(3, 1, out=d_out)
            cuda.cuda.synchronize()
            want_nco, _ = oracle.nco_words(want_plfg, 128, 16, 9, use_multiplier=True, enable_multiplying=bool(enable), multiplying_factor=factor,
                                           rounding_mode="Floor" if rounding == "RoundDown" else rounding)
            assert (from_dev_u32(d_nco) == want_nco).all(), (enable, hex(factor))
            fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
            assert (from_dev_u32(d_out) == oracle.chain_fixed_words(fft, mag, cfar, want_nco.reshape(3, 1024))).all()
    p.ncoParams.useMultiplier = False
    with R.RspChainVanilla(p) as ch:
        with pytest.raises(R.RspError):          # no register file without the multiplier
            ch.memWriteWord(p.ncoAddress.base, 1)


@pytest.mark.parametrize("log_n,Rc,G,lin", [(10, 32, 4, 1), (10, 8, 0, 0), (11, 16, 2, 1), (12, 32, 4, 1), (12, 31, 7, 0), (13, 5, 1, 1), (14, 32, 16, 1),
                                            (14, 24, 3, 0)])
def test_fast_os_kernel_bit_planes(cuda, oracle, log_n, Rc, G, lin):
    """chain_fixed_fast_os<S> (radix selection on bit planes) against the oracle's integer cfarOS core: every size, window sizes that are
    not powers of two, no guard cells, linear and log thresholds, full-scale random rows (negative magnitudes, ties), long runs of equal
    cells, ragged row counts, the detection list."""
    n = 1 << log_n
    rows = 9 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 2, n, seed=70 + log_n), synth.random_beats(2, n, seed=90 + log_n)])
    beats[0, n // 2:] = beats[0, 5]                      # equal cells: every candidate set keeps many members to the last plane
    p = fixed_params(n, 64, max(G, 1), ctype=R.OSGoldenCFARType)
    regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=lin, peakGrouping=0, cfarMode=0, windowCells=Rc, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast_os<{log_n}>", ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
        assert (got == want).all()
        assert k == int((want & 1).sum())
        assert set(map(tuple, dets[:, :2].tolist())) == set(map(tuple, np.argwhere((want & 1) == 1).tolist()))
        d_out.zero_()
        ch.run(d_in, 1, 1, out=d_out)                   # one row: the second chirp slot of a CTA (N = 1024) recomputes it
        cuda.cuda.synchronize()
        got1 = from_dev_u32(d_out)
        assert (got1[0] == want[0]).all() and not got1[1:].any()


@pytest.mark.parametrize("table_size", [100, 96, 3000, 128])
def test_nco_rasterized_mode(cuda, oracle, table_size):
    """FixedNCOParams(rasterizedMode = true) (NCOTable.scala:322-464): a table that is not a power of two (phaseWidth 9 and 14: the host-built
    beat table and the look-up in the kernel), the generator alone and the chain fed by it (fused first pass where the beat table exists),
    bit for bit against the oracle; a power-of-two table equals the standard mode."""
    p = R.rsp_chain_vanilla_app_params()
    p.ncoParams.rasterizedMode = True
    p.ncoParams.tableSize = table_size
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    order, repeats, start = [2, 0, 1, 3, 0], [2, 1, 3, 1, 2], 37
    total, n_rows = 6 * 1024 + 100, 6
    with R.RspChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        write_plfg(ch, p, oracle, start, CHIRPS, order, repeats, 0, 7)
        d_nco = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        d_plfg = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        ch.generate(total, d_nco, d_plfg)
        d_out = cuda.zeros((n_rows, 1024), dtype=cuda.int32, device="cuda")
        ch.run_from_generator(n_rows, 1, out=d_out)
        d_out2 = cuda.zeros((n_rows, 1024), dtype=cuda.int32, device="cuda")
        ch.run_from_generator_static(n_rows, 1, out=d_out2)
        cuda.cuda.synchronize()
        got_nco, got_plfg, got_out, got_out2 = from_dev_u32(d_nco), d_plfg.cpu().numpy(), from_dev_u32(d_out), from_dev_u32(d_out2)
    n = p.ncoParams
    assert n.phaseWidth == max(1, (table_size - 1).bit_length()) + 2
    want_plfg = oracle_plfg(oracle, p, start, CHIRPS, order, repeats, 0, 7).run(total)
    assert (got_plfg == want_plfg).all()
    want_nco, _ = oracle.nco_words(want_plfg, n.tableSize, n.tableWidth, n.phaseWidth, n.phaseAccEnable, rasterized=True)
    assert (got_nco == want_nco).all(), f"first difference at {np.argwhere(got_nco != want_nco)[:3].ravel()}"
    if table_size == 128:
        std, _ = oracle.nco_words(want_plfg, 128, 16, 9, True)
        assert (want_nco == std).all()
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want_out = oracle.chain_fixed_words(fft, mag, cfar, want_nco[:n_rows * 1024].reshape(n_rows, 1024))
    assert (got_out == want_out).all() and (got_out2 == want_out).all()


def test_unaligned_output_pointers_take_the_other_kernel_and_give_the_same_beats(cuda, oracle):
    """the bit-plane OS kernel stores 128-bit words and the fast kernels prefetch 16-byte aligned rows: a caller's buffer that starts at an
    odd word must still give the same beats (the planner falls back to 32-bit stores / the run-time-parameter kernel)"""
    n, log_n = 4096, 12
    beats = synth.chirps_fixed(5, n, seed=123)
    for ctype, regs in ((R.OSGoldenCFARType, dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)),
                        (GOSCFARType, dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)),
                        (CACFARType, dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=6, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4))):
        p = fixed_params(n, 64, 4, ctype=ctype)
        fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
        want = oracle.chain_fixed_words(fft, mag, cfar, beats)
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs)
            flat_in = cuda.zeros(5 * n + 3, dtype=cuda.int32, device="cuda")
            flat_out = cuda.zeros(5 * n + 3, dtype=cuda.int32, device="cuda")
            for off in (0, 1, 3):
                d_in = flat_in[off:off + 5 * n].view(5, n)
                d_in.copy_(to_dev(cuda, beats))
                d_out = flat_out[off:off + 5 * n].view(5, n)
                d_out.zero_()
                ch.run(d_in, 5, 1, out=d_out)
                d_f = flat_out.new_zeros(5 * n + 3)[off:off + 5 * n].view(5, n)
                ch.fft(d_in, 5, d_f)
                cuda.cuda.synchronize()
                assert (from_dev_u32(d_out) == want).all(), (ctype, off)
                assert (from_dev_u32(d_f) == oracle.fft_fixed_words(fft, beats)).all(), (ctype, off)


@pytest.mark.parametrize("log_n,W,G,gws,mode,lin", [(10, 16, 4, 4, 0, 1), (10, 64, 8, 8, 1, 0), (11, 32, 2, 4, 2, 1), (12, 64, 4, 4, 0, 1), (13, 16, 1, 1, 1, 1),
                                                    (14, 64, 8, 8, 0, 1), (14, 32, 3, 8, 2, 0)])
def test_fast_kernel_peak_grouping(cuda, oracle, log_n, W, G, gws, mode, lin):
    """peakGrouping = 1 on the specialised CA / GO / SO kernel (both neighbours come from the lagging guard, CFARCoreWithMem.scala:182-184;
    the left one only when guardCells == guardWindowSize): every size, linear / log, quiet and full-scale rows, ragged counts, detection list"""
    n = 1 << log_n
    rows = 7 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 2, n, seed=130 + log_n, noise=2.0 ** -3), synth.random_beats(2, n, seed=140 + log_n)])
    p = fixed_params(n, 64, gws)
    regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=lin, divSum=int(np.log2(W)), peakGrouping=1, cfarMode=mode, windowCells=W, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    regs0 = dict(regs, peakGrouping=0)
    _, _, cfar0 = oracle_cfgs(oracle, p, log_n, 2, regs0)
    want0 = oracle.chain_fixed_words(fft, mag, cfar0, beats)
    assert (want0 & 1).sum() >= (want & 1).sum() > 0         # the grouping can only remove peaks, and leaves some
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>", ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        k = int(det_count.item())
        assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
        assert k == int((want & 1).sum())
        assert set(map(tuple, det_list[:k].cpu().numpy()[:, :2].tolist())) == set(map(tuple, np.argwhere((want & 1) == 1).tolist()))


@pytest.mark.parametrize("log_n,W,G,gws,mode,lin", [(10, 16, 2, 2, 0, 1), (11, 32, 4, 4, 1, 1), (11, 32, 2, 4, 2, 0), (12, 64, 4, 4, 0, 1), (13, 16, 1, 1, 1, 1), (14, 64, 8, 8, 0, 1)])
def test_fast_f32_kernel_peak_grouping(cuda, oracle, log_n, W, G, gws, mode, lin):
    """peakGrouping = 1 on chain_f32_fast<S> (its own instantiation): thresholds 1e-4 against the fp64 dataflow, identical decided
    detections, guardCells equal to and below guardWindowSize (the left neighbour then reads as 0)"""
    n = 1 << log_n
    xc = synth.chirps_complex(4, n, seed=150 + log_n, noise=2.0 ** -4)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=n, logOrLinearMode=lin, divSum=int(np.log2(W)), peakGrouping=1, cfarMode=mode, windowCells=W, guardCells=G)
    scaler = 3.0 if lin else 1e-4
    thr, pk, variant = run_f32(cuda, f32_params(n, max(W, 16), gws), log_n, 2, regs, scaler, x32)
    assert variant == f"chain_f32_fast<{log_n}>", variant
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, scaler, W, G, guard_window_size=gws)
    check_f32(thr, pk, mag, want_thr, want_pk, f"fp32 pg mode {mode} lin {lin}")
    _, _, pk0 = f64_chain(oracle, xq, log_n, 2, dict(regs, peakGrouping=0), scaler, W, G, guard_window_size=gws)
    assert pk0.sum() >= want_pk.sum() > 0


@pytest.mark.parametrize("log_n,bp,W,G,mode,lin", [(10, 12, 64, 4, 0, 1), (10, 14, 16, 2, 1, 0), (11, 12, 32, 4, 2, 1), (12, 2, 64, 4, 0, 1), (13, 13, 16, 1, 1, 1),
                                                    (14, 12, 64, 8, 0, 1), (14, 14, 32, 3, 2, 0), (12, 15, 32, 2, 0, 1)])
def test_fast_kernel_binary_points(cuda, oracle, log_n, bp, W, G, mode, lin):
    """binPoint >= 2 on the specialised kernels (FftMagCfarChainVanillaApp / RxFftMagCfarTxChainApp have 12: src/main/scala/FftMagCfarChain.scala:76-120):
    the 3-multiplier twiddle product with per-product trim and the CFAR threshold with the operands' binary points, every size, quiet and
    full-scale rows (every wrap), the -32768 corners, ragged counts, the detection list, and the FFT block alone (fft_fixed_fast_kernel)"""
    n = 1 << log_n
    rows = 7 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 3, n, seed=160 + log_n, noise=2.0 ** -4), synth.random_beats(2, n, seed=170 + log_n),
                            np.full((1, n), 0x80008000, dtype=np.uint32)])
    p = fixed_params(n, 64, max(G, 1), bp=bp)
    p.cfarParams.protoScaler = FixedPoint(16, bp)          # the apps' choice: input, threshold and scaler share the binary point
    regs = dict(fftWin=n, thresholdScaler=int(3.25 * (1 << bp)) if lin else int(0.01 * (1 << bp)), logOrLinearMode=lin, divSum=int(np.log2(W)),
                peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, want_fft, _ = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>", ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        d_fft = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        ch.fft(d_in, rows, d_fft)
        cuda.cuda.synchronize()
        got, got_fft = from_dev_u32(d_out), from_dev_u32(d_fft)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
    assert (got_fft == want_fft).all(), f"FFT: {int((got_fft != want_fft).sum())} beats differ, first at {np.argwhere(got_fft != want_fft)[:3].tolist()}"
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    thr, bins, peak = decode_cfar_words(want, n)
    assert k == int(peak.sum()) and k > 0
    exp = {(r, b): int(thr[r, b]) for r, b in np.argwhere(peak == 1)}
    assert {(int(d[0]), int(d[1])): int(d[2]) & 0xFFFF for d in dets} == exp


@pytest.mark.parametrize("log_n,ib,tb,sb,mode,lin,fast", [(10, 0, 3, 6, 0, 1, True), (10, 0, 3, 6, 1, 0, True), (12, 4, 5, 2, 2, 1, True), (14, 0, 3, 6, 0, 1, True),
                                                           (11, 2, 0, 9, 1, 0, True), (10, 0, 8, 2, 0, 1, False)])
def test_fast_kernel_cfar_binary_points(cuda, oracle, log_n, ib, tb, sb, mode, lin, fast):
    """binary points in the CFAR alone (RspChainVanillaSpec: protoIn 16.0, protoThreshold 16.3, protoScaler 16.6; src/test/scala/RspChainVanillaTester.scala:229-231)
    behind a binPoint-0 transform: the specialised kernel's BPC instantiation, linear and log, other mixes; a threshold type finer than the
    context's binary point (the field would need a left shift) stays on the run-time-parameter kernel -- same beats either way"""
    n = 1 << log_n
    rows = 6 if log_n <= 12 else 3
    W, G = 32, 4
    beats = np.concatenate([synth.chirps_fixed(rows - 2, n, seed=180 + log_n, noise=2.0 ** -4), synth.random_beats(2, n, seed=190 + log_n)])
    p = fixed_params(n, 64, G)
    p.cfarParams.protoIn, p.cfarParams.protoThreshold, p.cfarParams.protoScaler = FixedPoint(16, ib), FixedPoint(16, tb), FixedPoint(16, sb)
    regs = dict(fftWin=n, thresholdScaler=int(3.5 * (1 << sb)) if lin else int(2.0 * (1 << sb)), logOrLinearMode=lin, divSum=5, peakGrouping=0, cfarMode=mode,
                windowCells=W, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert (ch.variant == f"chain_fixed_fast<{log_n}>") == fast, ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    thr, bins, peak = decode_cfar_words(want, n)
    assert k == int(peak.sum()) and k > 0
    assert {(int(d[0]), int(d[1])): int(d[2]) & 0xFFFF for d in dets} == {(r, b): int(thr[r, b]) for r, b in np.argwhere(peak == 1)}


@pytest.mark.parametrize("ctype,log_n,bp,W,G,kl,kr,lin", [("gos", 10, 12, 64, 4, 48, 48, 1), ("gos", 12, 12, 32, 2, 20, 7, 0), ("gos", 14, 14, 100, 3, 77, 77, 1),
                                                           ("os", 10, 12, 32, 4, 0, 0, 1), ("os", 12, 14, 24, 2, 0, 0, 0), ("os", 14, 12, 32, 8, 0, 0, 1),
                                                           ("gos", 11, 0, 16, 2, 12, 12, 1)])
def test_bit_plane_kernels_with_binary_points(cuda, oracle, ctype, log_n, bp, W, G, kl, kr, lin):
    """the OS / GOS bit-plane kernels behind a binPoint >= 2 transform (and, last case, with CFAR binary points behind a binPoint-0 one:
    input 0, threshold 3, scaler 6): bit-exact, specialised variant"""
    n = 1 << log_n
    rows = 6 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 2, n, seed=200 + log_n, noise=2.0 ** -4), synth.random_beats(2, n, seed=210 + log_n)])
    p = fixed_params(n, 128 if W > 64 else 64, max(G, 1), bp=bp, ctype=GOSCFARType if ctype == "gos" else R.OSGoldenCFARType)
    sb = bp
    if bp == 0:
        p.cfarParams.protoThreshold, sb = FixedPoint(16, 3), 6
    p.cfarParams.protoScaler = FixedPoint(16, sb)
    regs = dict(fftWin=n, thresholdScaler=int(2.5 * (1 << sb)) if lin else max(1, int(0.002 * (1 << sb))), logOrLinearMode=lin, peakGrouping=0, cfarMode=0, windowCells=W,
                guardCells=G, indexLagg=kl, indexLead=kr)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    got, variant = run_fixed(cuda, p, log_n, 2, regs, beats)
    assert variant == (f"chain_fixed_fast_gos<{log_n}>" if ctype == "gos" else f"chain_fixed_fast_os<{log_n}>"), variant
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    assert (want & 1).sum() > 0


@pytest.mark.parametrize("log_n,numpoints,bp,W,G,mode,lin,pg", [(6, 64, 0, 8, 1, 0, 1, 0), (6, 1024, 0, 16, 2, 1, 0, 0), (7, 128, 12, 16, 4, 2, 1, 0), (8, 1024, 0, 32, 4, 0, 1, 1),
                                                                 (8, 256, 14, 64, 8, 1, 1, 0), (9, 512, 0, 64, 4, 0, 0, 0), (9, 2048, 12, 32, 2, 2, 1, 0), (7, 512, 0, 32, 8, 0, 1, 1)])
def test_fast_kernel_below_1024_points(cuda, oracle, log_n, numpoints, bp, W, G, mode, lin, pg):
    """the specialised CA kernel at 64 ... 512 points (also as the run-time size of a larger generator): 4 ... 32 chirps per CTA, several chirps
    per warp, every bin on the edge path; binPoint 0 / 12 / 14, peak grouping, quiet and full-scale rows, row counts that do not fill the last CTA,
    the detection list, and the FFT block alone"""
    n = 1 << log_n
    rows = 75 if log_n < 12 else 11
    beats = np.concatenate([synth.chirps_fixed(rows - 9, n, seed=220 + log_n, noise=2.0 ** -4), synth.random_beats(8, n, seed=230 + log_n),
                            np.full((1, n), 0x80008000, dtype=np.uint32)])
    p = fixed_params(numpoints, 64, max(G, 1), bp=bp)
    p.cfarParams.protoScaler = FixedPoint(16, bp)
    regs = dict(fftWin=n, thresholdScaler=int(3.0 * (1 << bp)) if lin else max(1, int(0.004 * (1 << bp))), logOrLinearMode=lin, divSum=int(np.log2(W)),
                peakGrouping=pg, cfarMode=mode, windowCells=W, guardCells=G)
    if n <= 2 * W + 2 * G + 1:
        pytest.skip("fftWin must exceed both windows")
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, want_fft, _ = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>", ch.variant
        d_in = to_dev(cuda, beats)
        for k_rows in (rows, 1, min(33, rows - 1)):
            d_out = cuda.zeros_like(d_in)
            d_fft = cuda.zeros_like(d_in)
            det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
            det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
            ch.run(d_in, k_rows, 1, out=d_out, det_list=det_list, det_count=det_count)
            ch.fft(d_in, k_rows, d_fft)
            cuda.cuda.synchronize()
            got, got_fft = from_dev_u32(d_out), from_dev_u32(d_fft)
            assert (got_fft[:k_rows] == want_fft[:k_rows]).all() and not got_fft[k_rows:].any()
            assert (got[:k_rows] == want[:k_rows]).all(), f"{k_rows} rows: first difference at {np.argwhere(got[:k_rows] != want[:k_rows])[:3].tolist()}"
            assert not got[k_rows:].any()
            k = int(det_count.item())
            assert k == int((want[:k_rows] & 1).sum())
            assert set(map(tuple, det_list[:k].cpu().numpy()[:, :2].tolist())) == set(map(tuple, np.argwhere((want[:k_rows] & 1) == 1).tolist()))


def test_rsp_run_is_capturable_in_a_cuda_graph(cuda, oracle):
    """after one warm-up call (kernel attributes, occupancy query) rsp_run issues nothing but its kernel launch on the caller's stream, so a
    launch-bound loop of small batches can be captured once and replayed: the replayed graph gives the oracle's beats for new input contents"""
    n, log_n = 1024, 10
    p = fixed_params(n, 64, 4)
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    a, b = synth.chirps_fixed(4, n, seed=301), synth.chirps_fixed(4, n, seed=302)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        d_in = to_dev(cuda, a)
        d_out = cuda.zeros_like(d_in)
        ch.run(d_in, 4, 1, out=d_out)                      # warm-up outside the capture
        cuda.cuda.synchronize()
        g = cuda.cuda.CUDAGraph()
        s = cuda.cuda.Stream()
        with cuda.cuda.graph(g, stream=s):
            for _ in range(3):                             # three launches in the graph (the last one's result stays)
                ch.run(d_in, 4, 1, out=d_out, stream=s)
        d_out.zero_()
        g.replay()
        cuda.cuda.synchronize()
        assert (from_dev_u32(d_out) == oracle.chain_fixed_words(fft, mag, cfar, a)).all()
        d_in.copy_(to_dev(cuda, b))
        g.replay()
        cuda.cuda.synchronize()
        assert (from_dev_u32(d_out) == oracle.chain_fixed_words(fft, mag, cfar, b)).all()


@pytest.mark.parametrize("nco_kw,frame_num,inter", [(dict(), 0, 5), (dict(phaseAccEnable=False), 3, 9), (dict(useMultiplier=True, roundingMode="Convergent"), 0, 1),
                                                     (dict(rasterizedMode=True, tableSize=100), 2, 7), (dict(phaseWidth=12, tableSize=256), 0, 12)])
def test_generator_table_kernel_equals_the_closed_form(cuda, oracle, nco_kw, frame_num, inter):
    """rsp_generate without the PLFG words runs generator_table_kernel (the period's phase table + the beat table); with them the closed-form
    generator_kernel: same beats, for lengths that are not multiples of the run length, buffers that start at an odd word, several periods"""
    p = R.rsp_chain_vanilla_app_params()
    for k, v in nco_kw.items():
        setattr(p.ncoParams, k, v)
    order, repeats, start = [2, 0, 1, 3, 0], [2, 1, 3, 1, 2], 37
    frame_len = 2 * 2 + 320 + 3 * 475 + 213 + 2 * 320
    total = frame_num * frame_len + (frame_num - 1) * max(inter - 1, 0) if frame_num else 5 * frame_len + 77
    with R.RspChainVanilla(p) as ch:
        configure(ch, 10, 2, dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4))
        write_plfg(ch, p, oracle, start, CHIRPS, order, repeats, frame_num, inter)
        if nco_kw.get("useMultiplier"):
            ch.reg_write(R._abi.BLK_NCO, 0, 1); ch.reg_write(R._abi.BLK_NCO, 1, 0x3000)
        both = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        plfg = cuda.zeros(total, dtype=cuda.int32, device="cuda")
        ch.generate(total, both, plfg)
        for n_s, off in ((total, 0), (total - 3, 1), (5, 3), (1, 0)):
            flat = cuda.zeros(total + 4, dtype=cuda.int32, device="cuda")
            ch.generate(n_s, flat[off:off + total], None)
            cuda.cuda.synchronize()
            assert bool((flat[off:off + n_s] == both[:n_s]).all().item()), (n_s, off)
            assert not bool(flat[off + n_s:].any().item()) and not bool(flat[:off].any().item())


@pytest.mark.parametrize("ctype,log_n,W,G,kl,kr,lin", [("gos", 12, 64, 4, 48, 48, 1), ("gos", 10, 16, 2, 12, 5, 0), ("gos", 14, 100, 3, 77, 77, 1), ("os", 12, 32, 4, 0, 0, 1),
                                                       ("os", 11, 17, 0, 0, 0, 0)])
def test_cfar_block_alone_on_stored_profiles_with_an_ordered_statistic(cuda, oracle, ctype, log_n, W, G, kl, kr, lin):
    """BASELINE configs[3] "CFAR-only on stored 16-bit magnitude profiles" (SURVEY 8d): AXI4CFARBlock alone (rsp_cfar_run) with the hardware GOS
    core and with the golden cfarOS, exponential clutter + targets and full-range cells (negative ones too), against the oracle's CFAR"""
    n = 1 << log_n
    rows = 6 if log_n <= 12 else 3
    rng = np.random.default_rng(log_n * 7 + W)
    prof = np.floor(rng.exponential(256.0, (rows, n))).astype(np.int64)
    prof[:, [n // 8, n // 4, (3 * n) // 4]] += 4096
    prof[-1] = rng.integers(-32768, 32768, n)                 # every 16-bit value, negative cells included
    p = fixed_params(n, 128 if W > 64 else 64, max(G, 1), ctype=GOSCFARType if ctype == "gos" else R.OSGoldenCFARType)
    p.cfarParams.protoThreshold = FixedPoint(18, 0)
    regs = dict(fftWin=n, thresholdScaler=2 if lin else 300, logOrLinearMode=lin, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G, indexLagg=kl, indexLead=kr)
    _, _, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    words = (prof & 0xFFFF).astype(np.uint32)
    want = oracle.cfar_words(cfar, words)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        d_out = cuda.zeros((rows, n), dtype=cuda.int32, device="cuda")
        launches = ch.kernel_launches
        ch.cfar(to_dev(cuda, words), rows, out=d_out)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        assert ch.kernel_launches == launches + 1
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    assert (want & 1).sum() >= 3


def test_randomised_chain_configurations_against_the_oracle(cuda, oracle):
    """48 seeded random configurations over everything the specialised kernels and their fallbacks cover between them -- size (run-time sizes of a
    larger generator included), binary points (equal and mixed), CFAR core, windows, guard cells, thresholds, modes, peak grouping, row counts --
    each bit for bit against the oracle, whichever kernel the planner picks"""
    import os
    n_trials, seed = int(os.environ.get("RSP_FUZZ_TRIALS", "48")), int(os.environ.get("RSP_FUZZ_SEED", "20260"))   # a longer soak: RSP_FUZZ_TRIALS=600
    rng = np.random.default_rng(seed)
    picked = {}
    for trial in range(n_trials):
        log_n = int(rng.integers(6, 15))
        n = 1 << log_n
        extra = int(rng.choice([0, 0, 2])) if log_n <= 12 else 0          # generator larger than the transform: an even number of stages dropped
        numpoints = n << extra
        core = str(rng.choice(["ca", "ca", "ca", "gos", "os"])) if log_n >= 10 else "ca"
        bp = int(rng.choice([0, 0, 12, 14, 3]))
        mixed = bp == 0 and rng.random() < 0.3
        wmax = min(64, n // 4)
        W = int(rng.choice([w for w in (4, 8, 16, 32, 64) if w <= wmax])) if core != "gos" else int(rng.integers(3, min(100, n // 4)))
        if core == "os":
            W = min(W, 32)
        G = int(rng.integers(1, max(2, min(9, W))))
        lin = int(rng.integers(0, 2))
        mode = int(rng.integers(0, 3))
        pg = int(rng.random() < 0.2) if core == "ca" else 0
        ctype = {"ca": CACFARType, "gos": GOSCFARType, "os": R.OSGoldenCFARType}[core]
        p = fixed_params(numpoints, 128 if W > 64 else 64, max(G, 1) if rng.random() < 0.7 else 8, bp=bp, ctype=ctype)
        ib, tb, sb = (bp, bp, bp)
        if mixed:
            ib, tb, sb = 0, int(rng.integers(0, 5)), int(rng.integers(0, 8))
        p.cfarParams.protoIn, p.cfarParams.protoThreshold, p.cfarParams.protoScaler = FixedPoint(16, ib), FixedPoint(16, tb), FixedPoint(16, sb)
        scaler = int(rng.uniform(1.5, 4.0) * (1 << sb)) if lin else int(rng.uniform(0.0, 0.02) * (1 << sb))
        regs = dict(fftWin=n, thresholdScaler=scaler, logOrLinearMode=lin, divSum=max(0, int(np.log2(W))), peakGrouping=pg, cfarMode=mode, windowCells=W, guardCells=G,
                    indexLagg=int(rng.integers(1, W + 1)), indexLead=int(rng.integers(1, W + 1)))
        rows = int(rng.integers(1, 12)) if log_n >= 12 else int(rng.integers(1, 70))
        beats = np.concatenate([synth.chirps_fixed(rows, n, seed=1000 + trial, noise=2.0 ** -int(rng.integers(2, 8))), synth.random_beats(1, n, seed=2000 + trial)])[:rows] \
            if rows > 1 else synth.random_beats(1, n, seed=2000 + trial)
        try:
            fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
            want = oracle.chain_fixed_words(fft, mag, cfar, beats)
        except Exception:
            continue
        try:
            got, variant = run_fixed(cuda, p, log_n, 2, regs, beats)
        except R.RspError as e:          # refused configurations must be refused for a stated reason, never computed wrongly
            assert e.status in (R._abi.RSP_EINVAL, R._abi.RSP_EUNSUPPORTED), (trial, str(e))
            continue
        picked[variant.split("<")[0]] = picked.get(variant.split("<")[0], 0) + 1
        assert (got == want).all(), (trial, variant, dict(log_n=log_n, numpoints=numpoints, core=core, bp=bp, bps=(ib, tb, sb), W=W, G=G, lin=lin, mode=mode, pg=pg, rows=rows),
                                     np.argwhere(got != want)[:3].tolist())
    assert picked.get("chain_fixed_fast", 0) >= 10 and sum(picked.values()) >= 30, picked


def test_randomised_fp32_configurations_against_the_fp64_dataflow(cuda, oracle):
    """24 seeded random configurations of the DspReal flavour (sizes 256 ... 16384, window sizes that do and do not fit the fast kernel, guard cells,
    CA / GO / SO, linear / log, peak grouping): thresholds within 1e-4 of the fp64 dataflow, identical decided detections"""
    import os
    n_trials, seed = int(os.environ.get("RSP_FUZZ_TRIALS", "24")), int(os.environ.get("RSP_FUZZ_SEED", "99"))
    rng = np.random.default_rng(seed)
    fast = 0
    for trial in range(n_trials):
        log_n = int(rng.integers(8, 15))
        n = 1 << log_n
        W = int(rng.choice([w for w in (8, 16, 32, 48, 64, 128) if w <= n // 4]))
        G = int(rng.integers(1, min(9, W)))
        gws = G if rng.random() < 0.6 else 8
        mode, lin, pg = int(rng.integers(0, 3)), int(rng.integers(0, 2)), int(rng.random() < 0.3)
        rows = int(rng.integers(1, 5))
        xc = synth.chirps_complex(rows, n, seed=3000 + trial, noise=2.0 ** -int(rng.integers(3, 7)))
        x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
        xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
        regs = dict(fftWin=n, logOrLinearMode=lin, divSum=int(np.log2(W)) if W & (W - 1) == 0 else 5, peakGrouping=pg, cfarMode=mode, windowCells=W, guardCells=G)
        scaler = float(rng.uniform(2.0, 5.0)) if lin else float(rng.uniform(0.0, 2e-4))
        thr, pk, variant = run_f32(cuda, f32_params(n, max(W, 16) if W & (W - 1) == 0 else 64, gws), log_n, 2, regs, scaler, x32)
        fast += variant.startswith("chain_f32_fast")
        mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, scaler, W, G, guard_window_size=gws)
        scale = np.abs(want_thr).max()
        rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale)
        assert rel.max() < REL_TOL, (trial, variant, dict(log_n=log_n, W=W, G=G, gws=gws, mode=mode, lin=lin, pg=pg), float(rel.max()))
        sure = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale) > 10 * REL_TOL
        if pg:      # a neighbour within the tolerance of the cell under test leaves the local-maximum test undecided
            nb = np.minimum(np.abs(mag - np.roll(mag, -1, axis=1)), np.abs(mag - np.roll(mag, -G, axis=1))) / np.maximum(np.abs(mag), 1e-3 * scale) > 10 * REL_TOL
            sure &= nb
        assert (pk[sure] == want_pk[sure]).all(), (trial, variant, dict(log_n=log_n, W=W, G=G, gws=gws, mode=mode, lin=lin, pg=pg))
    assert fast >= n_trials // 4, fast


def test_randomised_generator_programmes_against_the_oracle(cuda, oracle):
    """32 seeded random PLFG programmes (tests/util.py:random_plfg_programme: 1-4 chirp types of 1-4 segments with random slopes, lengths, directions
    and resets, random chirp order and repeats, finite and endless frame counts, pauses 0 ... 63, starting points) through random NCO flavours, written
    through the register bus: PLFG words and NCO beats bit-exact from both generator kernels, the schedule's end refused one sample later, and the
    chain fed by the generator equal to the chain fed the oracle's beats.  The registers keep their Chisel widths (PLFGDspBlockMem.scala:49-53:
    segmentNums log2Ceil(maxNumOfSegments) = 2 bits, differentChirpsNum 3 bits), so 4 segments / 8 different chirps wrap to 0 as in the RTL: the
    oracle is given the wrapped values."""
    import os
    from tests.util import random_plfg_programme
    n_trials, seed = int(os.environ.get("RSP_FUZZ_TRIALS", "32")), int(os.environ.get("RSP_FUZZ_SEED", "4242"))
    rng = np.random.default_rng(seed)
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    for trial in range(n_trials):
        nco_kw, chirps, order, repeats, frame_num, inter, start, n_endless = random_plfg_programme(rng)
        p = R.rsp_chain_vanilla_app_params()
        for k, v in nco_kw.items():
            setattr(p.ncoParams, k, v)
        what = (trial, nco_kw, chirps, order, repeats, frame_num, inter, start)
        op = oracle_plfg(oracle, p, start, chirps, order, repeats, frame_num, inter)
        op.segment_nums = [len(s) & 3 for s in chirps]
        op.different_chirps_num = len(order) & 7
        want_plfg = op.run(n_endless if not frame_num else 1 << 18)
        total = len(want_plfg)
        assert total > 0 and (frame_num == 0 or total < 1 << 18), what
        n_rows = min(total // 1024, 4)
        with R.RspChainVanilla(p) as ch:
            configure(ch, 10, 2, regs)
            write_plfg(ch, p, oracle, start, chirps, order, repeats, frame_num, inter)
            d_nco = cuda.zeros(total, dtype=cuda.int32, device="cuda")
            d_plfg = cuda.zeros(total, dtype=cuda.int32, device="cuda")
            d_tab = cuda.zeros(total, dtype=cuda.int32, device="cuda")
            ch.generate(total, d_nco, d_plfg)       # generator_kernel (closed form)
            ch.generate(total, d_tab, None)         # generator_table_kernel
            if n_rows:
                d_out = cuda.zeros((n_rows, 1024), dtype=cuda.int32, device="cuda")
                ch.run_from_generator(n_rows, 1, out=d_out)
            cuda.cuda.synchronize()
            if frame_num:
                with pytest.raises(R.RspError):
                    ch.generate(total + 1, cuda.zeros(total + 1, dtype=cuda.int32, device="cuda"), None)
        got_plfg, got_nco, got_tab = d_plfg.cpu().numpy(), from_dev_u32(d_nco), from_dev_u32(d_tab)
        n = p.ncoParams                                 # (to_c has set the rasterized phase width by now)
        want_nco, _ = oracle.nco_words(want_plfg, n.tableSize, n.tableWidth, n.phaseWidth, n.phaseAccEnable, rasterized=n.rasterizedMode)
        assert (got_plfg == want_plfg).all(), (what, np.argwhere(got_plfg != want_plfg)[:3].ravel())
        assert (got_nco == want_nco).all(), (what, np.argwhere(got_nco != want_nco)[:3].ravel())
        assert (got_tab == want_nco).all(), (what, np.argwhere(got_tab != want_nco)[:3].ravel())
        if n_rows:
            fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
            assert (from_dev_u32(d_out) == oracle.chain_fixed_words(fft, mag, cfar, want_nco[:n_rows * 1024].reshape(n_rows, 1024))).all(), what


@pytest.mark.parametrize("log_n,bp,bps,W,G,gws,mode,lin", [(10, 12, None, 64, 4, 4, 0, 1),      # FftMagCfarChainVanillaApp's types with the testers' peakGrouping = 1
                                                            (10, 12, None, 32, 2, 4, 1, 0),      # guardCells below the generator's: the left neighbour reads 0
                                                            (12, 14, None, 16, 1, 1, 2, 1), (14, 12, None, 64, 8, 8, 0, 1), (13, 2, None, 32, 3, 8, 1, 1),
                                                            (10, 0, (0, 3, 6), 32, 4, 4, 0, 1),  # RspChainVanillaSpec's CFAR types behind a binPoint-0 transform
                                                            (14, 0, (0, 3, 6), 64, 4, 8, 2, 0), (11, 0, (4, 5, 2), 32, 4, 4, 1, 1),
                                                            (8, 12, None, 16, 2, 2, 0, 1)])       # below 1024 points
def test_fast_kernel_peak_grouping_with_binary_points(cuda, oracle, log_n, bp, bps, W, G, gws, mode, lin):
    """peakGrouping = 1 together with binary points (in the transform, in the CFAR, or both) on the specialised kernel's <PG, TW3 / BPC> instantiations:
    beats and detection list bit-exact, fewer peaks than without grouping, quiet / full-scale / corner rows"""
    n = 1 << log_n
    rows = 6 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 3, n, seed=260 + log_n, noise=2.0 ** -4), synth.random_beats(2, n, seed=270 + log_n),
                            np.full((1, n), 0x80008000, dtype=np.uint32)])
    p = fixed_params(n, 64, gws, bp=bp)
    if bps is None:
        p.cfarParams.protoScaler = FixedPoint(16, bp)
        sb = bp
    else:
        p.cfarParams.protoIn, p.cfarParams.protoThreshold, p.cfarParams.protoScaler = FixedPoint(16, bps[0]), FixedPoint(16, bps[1]), FixedPoint(16, bps[2])
        sb = bps[2]
    regs = dict(fftWin=n, thresholdScaler=int(3.25 * (1 << sb)) if lin else int(0.01 * (1 << sb)) + 1, logOrLinearMode=lin, divSum=int(np.log2(W)),
                peakGrouping=1, cfarMode=mode, windowCells=W, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats)
    fft0, mag0, cfar0 = oracle_cfgs(oracle, p, log_n, 2, dict(regs, peakGrouping=0))
    _, _, peak0 = decode_cfar_words(oracle.chain_fixed_words(fft0, mag0, cfar0, beats), n)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>", ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    thr, bins, peak = decode_cfar_words(want, n)
    assert k == int(peak.sum()) and 0 < k <= int(peak0.sum())
    assert {(int(d[0]), int(d[1])): int(d[2]) & 0xFFFF for d in dets} == {(r, b): int(thr[r, b]) for r, b in np.argwhere(peak == 1)}


def test_handles_on_two_streams_and_no_device_memory_left_behind(cuda, oracle):
    """two handles with different configurations interleaved on two streams (each call asynchronous on the stream it is given), results bit-exact; then
    150 handles created, used (chain, generator with its scratch, host-buffer run with its staging) and destroyed: the device's free memory returns to
    where it was (rsp_destroy frees tables, scratch and staging)"""
    n_a, n_b, rows = 1024, 4096, 64
    beats_a, beats_b = synth.chirps_fixed(rows, n_a, seed=301, noise=2.0 ** -4), synth.chirps_fixed(rows, n_b, seed=302, noise=2.0 ** -4)
    p_a, p_b = fixed_params(n_a, 64, 4), fixed_params(n_b, 64, 4, ctype=GOSCFARType)
    regs_a = dict(fftWin=n_a, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    regs_b = dict(fftWin=n_b, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
    want_a = oracle.chain_fixed_words(*oracle_cfgs(oracle, p_a, 10, 2, regs_a), beats_a)
    want_b = oracle.chain_fixed_words(*oracle_cfgs(oracle, p_b, 12, 2, regs_b), beats_b)
    s_a, s_b = cuda.cuda.Stream(), cuda.cuda.Stream()
    with R.FftMagCfarChainVanilla(p_a) as ch_a, R.FftMagCfarChainVanilla(p_b) as ch_b:
        configure(ch_a, 10, 2, regs_a); configure(ch_b, 12, 2, regs_b)
        d_a, d_b = to_dev(cuda, beats_a), to_dev(cuda, beats_b)
        outs_a, outs_b = [cuda.zeros_like(d_a) for _ in range(8)], [cuda.zeros_like(d_b) for _ in range(8)]
        cuda.cuda.synchronize()
        for i in range(8):
            ch_a.run(d_a, rows, 1, out=outs_a[i], stream=s_a)
            ch_b.run(d_b, rows, 1, out=outs_b[i], stream=s_b)
        s_a.synchronize(); s_b.synchronize()
        for i in range(8):
            assert (from_dev_u32(outs_a[i]) == want_a).all() and (from_dev_u32(outs_b[i]) == want_b).all(), i
    del d_a, d_b, outs_a, outs_b

    def one_handle(i):
        p = R.rsp_chain_vanilla_app_params()
        with R.RspChainVanilla(p) as ch:
            configure(ch, 10, 2, regs_a)
            write_plfg(ch, p, oracle, 37, CHIRPS, [2, 0, 1, 3, 0], [2, 1, 3, 1, 2], 0, 5)
            ch.run_from_generator(8 + i % 5, 1, out=scratch_out[:8 + i % 5])
            ch.run(scratch_in, 16, 1, out=scratch_out)
            if i % 10 == 0:
                ch.run_host(beats_a[:16], 16, 1)
            cuda.cuda.synchronize()
    scratch_in, scratch_out = to_dev(cuda, beats_a[:16]), cuda.zeros((16, n_a), dtype=cuda.int32, device="cuda")
    for i in range(10):      # warm-up: kernel attributes, the allocator's pools
        one_handle(i)
    cuda.cuda.synchronize()
    free0 = cuda.cuda.mem_get_info()[0]
    for i in range(150):
        one_handle(i)
    cuda.cuda.synchronize()
    free1 = cuda.cuda.mem_get_info()[0]
    assert free0 - free1 < (4 << 20), (free0, free1)     # (cudaMalloc works in 2 MiB pages)


def test_handles_from_four_host_threads(cuda, oracle):
    """four host threads, each with its own handle, stream and configuration (ctypes releases the GIL during the calls), 40 calls each: every result
    bit-exact -- the library's process-wide state (per-variant launch configuration, environment knobs, last-error text) is safe to share"""
    import threading
    cases = [(10, CACFARType, dict(thresholdScaler=4, logOrLinearMode=1, divSum=4, cfarMode=0, windowCells=16, guardCells=4)),
             (12, GOSCFARType, dict(thresholdScaler=2, logOrLinearMode=1, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)),
             (14, CACFARType, dict(thresholdScaler=4, logOrLinearMode=1, divSum=6, cfarMode=1, windowCells=64, guardCells=8)),
             (8, CACFARType, dict(thresholdScaler=3, logOrLinearMode=1, divSum=4, cfarMode=2, windowCells=16, guardCells=2))]
    jobs = []
    for i, (log_n, ctype, r) in enumerate(cases):
        n, rows = 1 << log_n, 12
        p = fixed_params(n, 64, 8 if log_n == 14 else 4, ctype=ctype)
        regs = dict(fftWin=n, peakGrouping=0, **r)
        beats = synth.chirps_fixed(rows, n, seed=400 + i, noise=2.0 ** -4)
        jobs.append((p, log_n, regs, beats, oracle.chain_fixed_words(*oracle_cfgs(oracle, p, log_n, 2, regs), beats)))
    errors = []

    def worker(p, log_n, regs, beats, want):
        try:
            cuda.cuda.set_device(0)
            s = cuda.cuda.Stream()
            with cuda.cuda.stream(s), R.FftMagCfarChainVanilla(p) as ch:
                configure(ch, log_n, 2, regs)
                d_in = to_dev(cuda, beats)
                outs = [cuda.zeros_like(d_in) for _ in range(4)]
                for it in range(40):
                    ch.run(d_in, beats.shape[0], 1, out=outs[it % 4], stream=s)
                    if it % 4 == 3:
                        s.synchronize()
                        for o in outs:
                            if not (from_dev_u32(o) == want).all():
                                errors.append((log_n, it, "beats differ"))
                                return
                            o.zero_()
        except Exception as e:      # noqa: BLE001 -- reported by the main thread
            errors.append((log_n, repr(e)))

    threads = [threading.Thread(target=worker, args=j) for j in jobs]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_randomised_whole_parameter_space_against_the_oracle(cuda, oracle):
    """60 seeded random configurations over the fixed-point chain's whole parameter space (tests/util.py:random_generic_configuration: sizes 8 ... 4096 and
    run-time sizes, binPoint 0 ... 14, all trim types, 3 / 4 multipliers, radix 2 / 2^2, DIF / DIT, keep-LSB stages, IFFT, every magnitude mode, the four
    CFAR cores with CASH, sendCut, peak grouping, linear / log) -- mostly the run-time-parameter kernel: bit-exact against the oracle, or refused"""
    import os
    from tests.util import random_generic_configuration
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "31337")))
    ran = refused = 0
    variants = set()
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "60"))):
        c = random_generic_configuration(rng, trial, fft_dir_reg=False, max_log=12)
        if c is None:
            continue
        p, log_n, sel, regs, beats, fft_dir, what = c
        try:
            with R.FftMagCfarChainVanilla(p) as ch:
                configure(ch, log_n, sel, regs)
                variants.add(ch.variant.split("<")[0])
                d_in = to_dev(cuda, beats)
                d_out = cuda.zeros_like(d_in)
                ch.run(d_in, beats.shape[0], 1, out=d_out)
                cuda.cuda.synchronize()
                got = from_dev_u32(d_out)
        except R.RspError as e:
            assert e.status in (R._abi.RSP_EINVAL, R._abi.RSP_EUNSUPPORTED), (what, str(e))
            refused += 1
            continue
        fft, mag, cfar = oracle_cfgs(oracle, p, log_n, sel, regs)
        want = oracle.chain_fixed_words(fft, mag, cfar, beats)
        assert (got == want).all(), (what, int((got != want).sum()), np.argwhere(got != want)[:3].tolist())
        ran += 1
    assert ran >= 30 and refused <= max(2, ran // 10), (ran, refused, variants)


def test_randomised_fp32_cores_against_the_fp64_dataflow(cuda, oracle):
    """30 seeded random configurations of the DspReal flavour over the four CFAR cores (tests/util.py:random_fp32_configuration: Mem, Lis / GOS, ASR / CASH,
    combined-window OS; window, guard, modes, ordered-statistic indices, linear / log, peak grouping; 64 ... 4096 points): thresholds within 1e-4 of the
    fp64 dataflow, decided detections identical"""
    import os
    from tests.util import check_fp32_against_f64, random_fp32_configuration
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "888")))
    variants = set()
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "30"))):
        c = random_fp32_configuration(rng, trial, sizes=(6, 8, 10, 11, 12))
        if c is None:
            continue
        p, log_n, regs, scaler, x32, what = c
        thr, pk, variant = run_f32(cuda, p, log_n, 2, regs, scaler, x32)
        variants.add(variant.split("<")[0])
        check_fp32_against_f64(oracle, what, log_n, regs, scaler, x32, thr, pk, tol=REL_TOL)
    assert "chain_f32" in variants, variants


def test_randomised_wide_fft_blocks(cuda, oracle):
    """30 seeded random configurations of the FFT block with wide beats (fft_wide_kernel): 16 ... 4096 points and run-time sizes, 17 ... 32-bit data,
    9 ... 32-bit twiddles, any binary point, random grow stages and keep-LSB stages, every trim type, 3 / 4 multipliers, natural and bit-reversed output;
    beats and the overflow status bit-exact against the oracle, or refused"""
    import os
    from tests.util import TRIM_TYPES
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "1234")))
    ran = refused = 0
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "30"))):
        s = int(rng.integers(4, 13))
        n = 1 << s
        dw = int(rng.integers(17, 33)) if rng.random() < 0.8 else 16
        grow = [int(rng.random() < 0.35) for _ in range(s)] if (dw == 16 or rng.random() < 0.5) else [0] * s
        if dw == 16 and not any(grow):
            grow[int(rng.integers(0, s))] = 1           # (16 bits without grow is the 32-bit-beat kernels' configuration)
        if dw + sum(grow) > 32 + 0 and dw > 16:
            grow = [0] * s if dw + sum(grow) > 40 else grow
        tw = int(rng.integers(9, min(32, dw + 8) + 1))
        bp = int(rng.integers(0, dw - 1))
        keep = [bool(rng.random() < 0.85) for _ in range(s)]
        run_time = bool(rng.random() < 0.3)
        log_n = s - 2 * int(rng.integers(0, (s - 2) // 2 + 1)) if run_time else s
        kw = dict(trimType=str(rng.choice(TRIM_TYPES)), use4Muls=bool(rng.integers(0, 2)), useBitReverse=bool(rng.random() < 0.7), runTime=run_time)
        what = (trial, dict(n=n, log_n=log_n, dw=dw, bp=bp, tw=tw, grow=grow, keep=keep, **kw))
        m, rows = 1 << log_n, 3
        re = rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), (rows, m)); im = rng.integers(-(1 << (dw - 1)), 1 << (dw - 1), (rows, m))
        re[1] >>= 3; im[1] >>= 3
        beats = pack_beats64(re, im, dw)
        try:
            p = wide_params(n, dw, bp, tw, grow, keep, overflowReg=True, **kw)
            with R.FftMagCfarChainVanilla(p) as ch:
                if run_time:
                    ch.memWriteWord(p.fftAddress.base, log_n)
                k = ch.fft_out_beat_bytes // 8
                d_in = cuda.from_numpy(beats.view(np.int64)).cuda()
                d_out = cuda.zeros((rows, m, k), dtype=cuda.int64, device="cuda")
                ch.fft(d_in, rows, d_out)
                got = d_out.cpu().numpy().view(np.uint64)
                ovf = ch.fft_overflow_flags()
        except R.RspError as e:
            assert e.status in (R._abi.RSP_EINVAL, R._abi.RSP_EUNSUPPORTED), (what, str(e))
            refused += 1
            continue
        fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, dict(fftWin=max(m, 64)))
        want, want_ovf = oracle.fft_fixed_words64(fft, beats)
        assert (got.reshape(want.shape) == want).all(), (what, np.argwhere(got.reshape(want.shape) != want)[:3].tolist())
        assert ovf == want_ovf << (s - log_n), (what, hex(ovf), hex(want_ovf))
        ran += 1
    print(f"wide FFT: {ran} ran, {refused} refused")
    assert ran >= 12 and refused <= ran, (ran, refused)


@pytest.mark.parametrize("ctype,log_n,bp,W,G,gws,kl,kr,lin", [("gos", 12, 0, 64, 4, 4, 48, 48, 1),      # C4 reading (A) with the testers' peakGrouping = 1
                                                               ("gos", 10, 0, 16, 2, 4, 12, 5, 0),      # guardCells below the generator's: x[n + 1] reads 0
                                                               ("gos", 14, 12, 100, 3, 3, 77, 77, 1), ("gos", 11, 0, 24, 1, 1, 1, 24, 1),
                                                               ("gos", 13, 14, 40, 2, 2, 30, 9, 0)])
def test_bit_plane_kernels_peak_grouping(cuda, oracle, ctype, log_n, bp, W, G, gws, kl, kr, lin):
    """peakGrouping = 1 on the GOS bit-plane kernel (the cfarOS golden has no peak grouping: refused; neighbours x[n + 1] and x[n - 1] from the guard registers, CFARCoreWithLis.scala:191-192): the chain
    and the CFAR block alone on stored profiles, bit-exact on the specialised variant, fewer peaks than without grouping, the detection list"""
    n = 1 << log_n
    rows = 6 if log_n <= 12 else 3
    beats = np.concatenate([synth.chirps_fixed(rows - 2, n, seed=280 + log_n, noise=2.0 ** -4), synth.random_beats(2, n, seed=290 + log_n)])
    p = fixed_params(n, 128 if W > 64 else 64, gws, bp=bp, ctype=GOSCFARType if ctype == "gos" else R.OSGoldenCFARType)
    p.cfarParams.protoScaler = FixedPoint(16, bp)
    regs = dict(fftWin=n, thresholdScaler=int(2.5 * (1 << bp)) if lin else max(1, int(0.002 * (1 << bp))), logOrLinearMode=lin, peakGrouping=1, cfarMode=0, windowCells=W,
                guardCells=G, indexLagg=kl, indexLead=kr)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, _, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    _, _, peak0 = decode_cfar_words(oracle.chain_fixed_words(*oracle_cfgs(oracle, p, log_n, 2, dict(regs, peakGrouping=0)), beats), n)
    name = f"chain_fixed_fast_gos<{log_n}>" if ctype == "gos" else f"chain_fixed_fast_os<{log_n}>"
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == name, ch.variant
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        d_out2 = cuda.zeros_like(d_in)
        launches = ch.kernel_launches
        ch.cfar(to_dev(cuda, np.ascontiguousarray(want_mag).astype(np.uint32)), rows, out=d_out2)      # the CFAR block alone on the oracle's magnitudes
        cuda.cuda.synchronize()
        assert ch.kernel_launches == launches + 1
        got, got2 = from_dev_u32(d_out), from_dev_u32(d_out2)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
    assert (got2 == want).all(), f"CFAR alone: {int((got2 != want).sum())} beats differ, first at {np.argwhere(got2 != want)[:3].tolist()}"
    thr, bins, peak = decode_cfar_words(want, n)
    assert k == int(peak.sum()) and 0 < k <= int(peak0.sum())
    assert set((int(d[0]), int(d[1])) for d in dets) == set(map(tuple, np.argwhere(peak == 1).tolist()))


@pytest.mark.parametrize("log_n,bp", [(10, 0), (14, 0), (12, 1), (6, 0), (10, 12), (14, 14), (11, 2), (13, 15), (8, 12)])
def test_convergent_rounding_on_the_compile_time_modes(cuda, oracle, log_n, bp):
    """trimType = Convergent (what the reference's SQNR specs elaborate, Radix22SQNRSpec.scala) on the run-time-parameter kernel's compile-time modes
    <fast-convergent> / <fast3-convergent> (before: the <generic> mode, 7 Gsamples/s): chain and FFT block alone bit-exact on full-range random
    beats (every tie, every wrap), chirps and the -32768 corners"""
    n = 1 << log_n
    rows = 7 if log_n <= 12 else 4
    beats = np.concatenate([synth.random_beats(rows - 3, n, seed=310 + log_n), synth.chirps_fixed(2, n, seed=320 + log_n, noise=2.0 ** -4),
                            np.full((1, n), 0x80008000, dtype=np.uint32)])
    W, G = (16, 2) if n >= 128 else (4, 1)
    p = fixed_params(n, 64 if n >= 256 else 4, G, bp=bp)
    p.fftParams.trimType = "Convergent"
    p.cfarParams.protoScaler = FixedPoint(16, bp)
    regs = dict(fftWin=n, thresholdScaler=int(3.0 * (1 << bp)), logOrLinearMode=1, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want, want_fft, _ = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed<{log_n},{'fast-convergent' if bp <= 1 else 'fast3-convergent'}>", ch.variant
        d_in = to_dev(cuda, beats)
        d_out, d_fft = cuda.zeros_like(d_in), cuda.zeros_like(d_in)
        ch.run(d_in, rows, 1, out=d_out)
        ch.fft(d_in, rows, d_fft)
        cuda.cuda.synchronize()
        got, got_fft = from_dev_u32(d_out), from_dev_u32(d_fft)
    assert (got_fft == want_fft).all(), f"FFT: {int((got_fft != want_fft).sum())} beats differ"
    assert (got == want).all(), f"{int((got != want).sum())} beats differ, first at {np.argwhere(got != want)[:3].tolist()}"
