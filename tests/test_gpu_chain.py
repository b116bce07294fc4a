"""Parity of the CUDA path (through the C ABI) against the CPU oracle, fixed-point chain.  Bit-exact."""
import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import synth
from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint,
                                    GOSCACFARType, GOSCFARType, MAGParams)
from tests.util import configure, decode_cfar_words, from_dev_u32, oracle_cfgs, to_dev

pytestmark = pytest.mark.gpu


def chain_params(n=1024, bp=0, trim="RoundHalfUp", cfar_type=CACFARType, cash=False, W=64, G=4, **fft_kw):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=bp,
                                  trimType=trim, **fft_kw),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=bp, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, bp), protoThreshold=FixedPoint(16, bp), protoScaler=FixedPoint(16, bp),
                              leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, includeCASH=cash,
                              minSubWindowSize=4 if cash else None, CFARAlgorithm=cfar_type))


def run_and_compare(cuda, oracle, params, log_n, sel, regs, beats):
    n = 1 << log_n
    rows = beats.shape[0]
    with R.FftMagCfarChainVanilla(params) as ch:
        configure(ch, log_n, sel, regs)
        d_in = to_dev(cuda, beats)
        d_out = cuda.empty((rows, n), dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        variant = ch.variant
    fft, mag, cfar = oracle_cfgs(oracle, params, log_n, sel, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats, n_threads=oracle.max_threads())
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"{variant}: {bad.shape[0]} beats differ, first at {bad[0]}: got {got[tuple(bad[0])]:#x} want {want[tuple(bad[0])]:#x}"
    return got


def test_c2_frame_ca(cuda, oracle):
    """BASELINE config 1: one frame 256 x 1024, fixed bp 0, JPL, CA-CFAR guard 4 ref 16"""
    p = chain_params(1024, 0)
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    beats = synth.chirps_fixed(256, 1024)
    got = run_and_compare(cuda, oracle, p, 10, 2, regs, beats)
    thr, bins, peak = decode_cfar_words(got, 1024)
    assert (bins == np.arange(1024)[None, :]).all()
    # the two synthetic targets of row 0 are detected
    assert peak[0, 128] == 1 and peak[0, 256] == 1


@pytest.mark.parametrize("log_n", [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14])
def test_chain_all_sizes_random(cuda, oracle, log_n):
    """every transform size, full-range random beats (all overflow corners), CA/GO/SO"""
    n = 1 << log_n
    W = max(1, min(16, n // 8))
    G = 1
    p = chain_params(n, 0, W=max(2, W), G=max(1, G))
    beats = synth.random_beats(37 if log_n < 13 else 5, n, seed=log_n)
    if not (n > 2 * W + 2 * G + 1):
        # 2 and 4 points: no CFAR window fits (fftWin > 2 W + 2 G + 1, CFARCoreWithMem.scala:31-32): the FFT block alone, and the chain refuses
        fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, dict(fftWin=n))
        want = oracle.fft_fixed_words(fft, beats)
        with R.FftMagCfarChainVanilla(p) as ch:
            ch.memWriteWord(p.fftAddress.base, log_n)
            d_out = cuda.empty((beats.shape[0], n), dtype=cuda.int32, device="cuda")
            ch.fft(to_dev(cuda, beats), beats.shape[0], d_out)
            assert (from_dev_u32(d_out) == want).all()
            configure(ch, log_n, 2, dict(fftWin=n, windowCells=W, guardCells=G))
            with pytest.raises(R.RspError):
                ch.run(to_dev(cuda, beats), beats.shape[0], 1, out=d_out)
        return
    for mode in (0, 1, 2):
        regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=max(0, W.bit_length() - 1), peakGrouping=mode & 1,
                    cfarMode=mode, windowCells=W, guardCells=G)
        run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)


@pytest.mark.parametrize("bp,trim,use4", [(12, "RoundHalfUp", False), (14, "RoundHalfUp", False), (12, "Convergent", False),
                                           (12, "RoundHalfUp", True), (0, "Convergent", False), (12, "RoundDown", False)])
def test_chain_binpoints_and_trims(cuda, oracle, bp, trim, use4):
    p = chain_params(1024, bp, trim, use4Muls=use4)
    regs = dict(fftWin=1024, thresholdScaler=int(3.5 * (1 << bp)), logOrLinearMode=1, divSum=5, cfarMode=1, windowCells=32, guardCells=4)
    beats = np.concatenate([synth.chirps_fixed(24, 1024, seed=5), synth.random_beats(8, 1024, seed=6)])
    for sel in (1, 2, 3):
        run_and_compare(cuda, oracle, p, 10, sel, regs, beats)


def test_runtime_size_smaller_than_numpoints(cuda, oracle):
    p = chain_params(1024, 0)
    for log_n in (6, 8):
        n = 1 << log_n
        regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=2, cfarMode=0, windowCells=4, guardCells=2)
        run_and_compare(cuda, oracle, p, log_n, 2, regs, synth.random_beats(19, n, seed=log_n, full_range=False))
    with R.FftMagCfarChainVanilla(p) as ch:
        ch.memWriteWord(p.fftAddress.base, 9)   # odd number of dropped stages: not a radix-2^2 size
        ch.memWriteWord(p.cfarAddress.base, 512)
        d = cuda.zeros((1, 512), dtype=cuda.int32, device="cuda")
        with pytest.raises(R.RspError):
            ch.run(d, 1, 1, out=d.clone())


def test_log_mode_and_threshold_field(cuda, oracle):
    p = chain_params(256, 0, W=16, G=2)
    regs = dict(fftWin=256, thresholdScaler=40, logOrLinearMode=0, divSum=4, cfarMode=0, windowCells=16, guardCells=2)
    run_and_compare(cuda, oracle, p, 8, 3, regs, synth.chirps_fixed(33, 256, seed=9))


@pytest.mark.parametrize("ctype,cash", [(GOSCFARType, False), (GOSCACFARType, False), (CACFARType, True)])
def test_cfar_cores(cuda, oracle, ctype, cash):
    """Lis core (GOS, both algorithms of GOSCA) and ASR core (CASH) behind the same chain"""
    p = chain_params(512, 0, cfar_type=ctype, cash=cash, W=32, G=4)
    beats = synth.chirps_fixed(21, 512, seed=3)
    for mode in (0, 1, 2, 3):
        for alg in ((0, 1) if ctype == GOSCACFARType else (0,)):
            regs = dict(fftWin=512, thresholdScaler=3, logOrLinearMode=1, divSum=4, peakGrouping=1, cfarAlgorithm=alg, cfarMode=mode,
                        windowCells=16, guardCells=3, indexLagg=12, indexLead=5, subWindowSize=4)
            run_and_compare(cuda, oracle, p, 9, 2, regs, beats)


def test_os_cfar_c4(cuda, oracle):
    """BASELINE config 3 reading (A): GOS, 64 cells per side, k = 48, 4096-point profiles"""
    p = chain_params(4096, 0, cfar_type=GOSCFARType, W=64, G=4)
    regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4,
                indexLagg=48, indexLead=48)
    run_and_compare(cuda, oracle, p, 12, 2, regs, synth.chirps_fixed(16, 4096, seed=4))


def test_blocks_alone(cuda, oracle):
    """AXI4FFTBlock, AXI4LogMagMuxBlock, AXI4CFARBlock used on their own equal the oracle's stages"""
    p = chain_params(1024, 12)
    regs = dict(fftWin=1024, thresholdScaler=int(3.5 * 4096), logOrLinearMode=1, divSum=5, cfarMode=0, windowCells=32, guardCells=4)
    beats = synth.chirps_fixed(16, 1024, seed=8)
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want_out, want_fft, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, want_stages=True)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        d_in = to_dev(cuda, beats)
        d_f = cuda.empty_like(d_in); d_m = cuda.empty_like(d_in); d_c = cuda.empty_like(d_in)
        ch.fft(d_in, 16, d_f)
        ch.mag(d_f, 16 * 1024, d_m)
        ch.cfar(d_m, 16, out=d_c)
        cuda.cuda.synchronize()
        assert (from_dev_u32(d_f) == want_fft).all()
        assert (from_dev_u32(d_m) == want_mag).all()
        assert (from_dev_u32(d_c) == want_out).all()
        assert ch.kernel_launches == 3


def test_run_host_and_detections(cuda, oracle):
    p = chain_params(1024, 0)
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=4, cfarMode=0, windowCells=16, guardCells=4)
    rows = 4500   # several pipeline chunks with a ragged tail
    beats = np.tile(synth.chirps_fixed(250, 1024, seed=2), (18, 1))[:rows]
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, beats[:250], n_threads=oracle.max_threads())
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        out, _, dets, n_dets = ch.run_host(beats, rows, 1, max_dets=1 << 20)
    assert (out[:250] == want).all()
    assert (out[250 * 17:] == want[:rows - 250 * 17]).all()
    _, bins, peak = decode_cfar_words(out, 1024)
    assert n_dets == int(peak.sum()) == dets.shape[0]
    got = set(map(tuple, dets[:, :2].tolist()))
    exp = set(map(tuple, np.argwhere(peak == 1).tolist()))
    assert got == exp


def test_invalid_parameters_raise(cuda):
    p = chain_params(1024, 0)
    p.cfarParams.guardWindowSize = 0   # CFARParams.scala:41
    with pytest.raises(R.RspError) as e:
        R.FftMagCfarChainVanilla(p)
    assert e.value.status == R._abi.RSP_EINVAL
    p = chain_params(1024, 0)
    with R.FftMagCfarChainVanilla(p) as ch:
        ch.memWriteWord(p.cfarAddress.base, 1024)
        ch.memWriteWord(p.cfarAddress.base + 7 * 4, 64)
        ch.memWriteWord(p.cfarAddress.base + 8 * 4, 4)
        ch.memWriteWord(p.fftAddress.base, 6)   # fftWin (1024) != transform size (64)
        d = cuda.zeros((1, 64), dtype=cuda.int32, device="cuda")
        with pytest.raises(R.RspError):
            ch.run(d, 1, 1, out=d.clone())


# ------------------------------------------------------------------------------------------- the specialised fast kernel
def _fast_params(n, W, G, thr_w=16):
    p = chain_params(n, 0, W=W, G=G)
    p.cfarParams.protoThreshold = FixedPoint(thr_w, 0)
    return p


@pytest.mark.parametrize("log_n,W,G", [(10, 16, 4), (10, 64, 4), (10, 32, 8), (11, 32, 4), (12, 64, 4), (13, 16, 2), (14, 64, 8)])
def test_fast_kernel_variants(cuda, oracle, log_n, W, G):
    """chain_fixed_fast<S>: CA/GO/SO x linear/log, ragged row counts (two chirps per CTA), edge iterations (W + G > 64 at
    N = 1024), threshold fields of other widths, full-range random rows next to realistic chirps"""
    n = 1 << log_n
    rows = 37 if log_n <= 12 else 9
    beats = np.concatenate([synth.chirps_fixed(rows - 5, n, seed=log_n), synth.random_beats(5, n, seed=100 + log_n)])
    for mode, lin, thr_w, scaler in ((0, 1, 16, 4), (1, 1, 16, 3), (2, 0, 16, 700), (0, 1, 12, 5), (2, 1, 20, -3), (3, 1, 16, 2)):
        p = _fast_params(n, W, G, thr_w)
        regs = dict(fftWin=n, thresholdScaler=scaler & 0xFFFF, logOrLinearMode=lin, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=mode,
                    windowCells=W, guardCells=G)
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs)
            assert ch.variant == f"chain_fixed_fast<{log_n}>", ch.variant
        run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)


def test_fast_kernel_detection_list(cuda, oracle):
    """the out-of-line detection path of the fast kernel reports exactly the beats with peak = 1 (chirp, bin, threshold, cut)"""
    p = _fast_params(1024, 16, 4)
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    rows = 301
    beats = synth.chirps_fixed(rows, 1024, seed=21)
    fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
    want, _, want_mag = oracle.chain_fixed_words(fft, mag, cfar, beats, n_threads=oracle.max_threads(), want_stages=True)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        d_in = to_dev(cuda, beats)
        d_out = cuda.empty((rows, 1024), dtype=cuda.int32, device="cuda")
        det_list = cuda.zeros((1 << 16, 4), dtype=cuda.int32, device="cuda")
        det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
        ch.run(d_in, rows, 1, out=d_out, det_list=det_list, det_count=det_count)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
        k = int(det_count.item())
        dets = det_list[:k].cpu().numpy()
    assert (got == want).all()
    thr, bins, peak = decode_cfar_words(want, 1024)
    exp = {(r, b): (int(thr[r, b]), int(np.int16(want_mag[r, b] & 0xFFFF))) for r, b in np.argwhere(peak == 1)}
    assert k == len(exp)
    assert {(int(d[0]), int(d[1])): (int(d[2]) & 0xFFFF, int(d[3])) for d in dets} == exp


def test_c5_full_size_rows_are_independent(cuda, oracle):
    """BASELINE config 4 at full size: 8 RX x 4096 chirps x 16384 points (2^29 samples, 2 GiB in + 2 GiB out).  The oracle
    checks 8 distinct rows; the size-independent property is that a row's result depends on that row alone: every replica of
    the 8-row block must equal the first, wherever it sits in the batch."""
    n, log_n, W, G = 16384, 14, 64, 8
    p = _fast_params(n, W, G)
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=6, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    base = np.concatenate([synth.chirps_fixed(7, n, seed=77), synth.random_beats(1, n, seed=78)])
    n_rx, n_chirps = 8, 4096
    rows = n_rx * n_chirps
    fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.chain_fixed_words(fft, mag, cfar, base, n_threads=oracle.max_threads())
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        d_in = to_dev(cuda, base).repeat(rows // 8, 1).contiguous()
        d_out = cuda.empty_like(d_in)
        ch.run(d_in, n_chirps, n_rx, out=d_out)
        cuda.cuda.synchronize()
        assert ch.variant == "chain_fixed_fast<14>"
        first = d_out[:8]
        assert (from_dev_u32(first) == want).all()
        assert bool((d_out.view(rows // 8, 8, n) == first.unsqueeze(0)).all().item())


@pytest.mark.parametrize("log_n,W,G,kl,kr", [(10, 16, 2, 12, 12), (10, 64, 4, 48, 48), (11, 32, 0, 24, 7), (12, 64, 4, 48, 20), (13, 8, 1, 1, 8),
                                              (14, 128, 8, 96, 96), (11, 24, 2, 20, 20), (12, 40, 3, 1, 40), (10, 72, 1, 72, 1),
                                              (11, 13, 2, 5, 13), (12, 100, 3, 77, 30), (14, 97, 0, 50, 50), (13, 3, 0, 2, 2)])
def test_fast_gos_kernel_variants(cuda, oracle, log_n, W, G, kl, kr):
    """chain_fixed_fast_gos<S> (rank-counting OS-CFAR): CA/GO/SO combination of the two order statistics, linear and log threshold,
    equal and different indexLagg / indexLead, ragged row counts, full-range random rows (ties, negative magnitudes)"""
    n = 1 << log_n
    rows = 19 if log_n <= 12 else 5
    beats = np.concatenate([synth.chirps_fixed(rows - 3, n, seed=log_n), synth.random_beats(3, n, seed=200 + log_n)])
    beats[-1, n // 2:] = beats[-1, 7]          # long runs of equal cells
    # the packed rank counting takes rows whose cells all lie in 0 .. 32766, the scalar one the others: a row at each side of the limit
    # (DC rows: bin 0 = the beat itself, JPL magnitude 32766 / 32767 / wrapped negative), and a loud row next to a quiet one
    def const_row(re, im): return np.full(n, ((re & 0xFFFF) << 16) | (im & 0xFFFF), dtype=np.uint32)
    beats = np.concatenate([beats, np.stack([const_row(32766, 0), const_row(32767, 0), const_row(32767, 32767), const_row(0, -32768)]),
                            synth.chirps_fixed(2, n, seed=300 + log_n)])
    p = chain_params(n, 0, cfar_type=GOSCFARType, W=max(1 << (W - 1).bit_length(), 16), G=max(G, 1))
    for mode, lin, scaler in ((0, 1, 2), (1, 1, 3), (2, 0, 500), (3, 1, -2)):
        regs = dict(fftWin=n, thresholdScaler=scaler & 0xFFFF, logOrLinearMode=lin, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G,
                    indexLagg=kl, indexLead=kr)
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs)
            assert ch.variant == f"chain_fixed_fast_gos<{log_n}>", ch.variant
        run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)


@pytest.mark.parametrize("log_n", [10, 12, 14])
def test_fast_kernel_quiet_and_loud_rows(cuda, oracle, log_n):
    """the fast kernel leaves the 16-bit wraps of the butterflies out for rows whose beats are all within +-23150 (they cannot wrap)
    and keeps them for the others, warp by warp in the first pass and chirp by chirp afterwards: rows at the limit, one beat over it,
    full-scale square waves (the -32768 / +32767 corners that do wrap), and a loud beat seen by only one warp of the chirp"""
    n = 1 << log_n
    rng = np.random.default_rng(log_n)
    rows = []
    for amp in (23150, 23151, 32767):
        re = rng.integers(-amp, amp + 1, n); im = rng.integers(-amp, amp + 1, n)
        re[::7] = amp; im[::5] = -amp
        rows.append((re, im))
    sq = np.where((np.arange(n) // 3) % 2 == 0, 32767, -32768)
    rows.append((sq, -sq - 1))
    rows.append((np.full(n, -32768), np.full(n, -32768)))
    for pos in (0, n // 2 + 5, n - 1):          # quiet noise with a single loud beat
        re = rng.integers(-2000, 2001, n); im = rng.integers(-2000, 2001, n)
        re[pos] = -32768
        rows.append((re, im))
    beats = np.stack([((r.astype(np.int64) & 0xFFFF) << 16 | (i.astype(np.int64) & 0xFFFF)).astype(np.uint32) for r, i in rows])
    beats = np.concatenate([beats, synth.chirps_fixed(3, n, seed=77)])
    p = _fast_params(n, 16, 4)
    regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        assert ch.variant == f"chain_fixed_fast<{log_n}>"
    run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)


def test_window_functions(cuda, oracle):
    """FFTParams.windowFunc (WindowFunc.scala, SDFFFT.scala:53-66,233-234): every window type in front of the forward transform, a
    narrower coefficient type, a run-time size below numPoints (first n ROM entries), no window on the IFFT; whole chain and FFT alone"""
    from oracle.glue import apply_window
    from rsp_chains_b200 import WindowFunctionTypes as WT
    n = 1024
    rng = np.random.default_rng(5)
    user = WT.User(14, dv=list(rng.uniform(-1.5, 1.9, n)))
    beats = np.concatenate([synth.chirps_fixed(6, n, seed=21), synth.random_beats(3, n, seed=22)])
    regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    for win in (WT.Hanning(), WT.Hamming(16, 0.54, 0.46), WT.Blackman(12), WT.Triangular(10), user):
        p = chain_params(n, 0, W=16, G=4)
        p.fftParams.windowFunc = win
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, 10, 2, regs)
            assert ch.variant == "chain_fixed_fast<10>", ch.variant
        # the product applies the window itself: compare it on the raw beats with the oracle on the windowed beats
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, 10, 2, regs)
            d_out = cuda.empty((beats.shape[0], n), dtype=cuda.int32, device="cuda")
            ch.run(to_dev(cuda, beats), beats.shape[0], 1, out=d_out)
            d_fft = cuda.empty((beats.shape[0], n), dtype=cuda.int32, device="cuda")
            ch.fft(to_dev(cuda, beats), beats.shape[0], d_fft)
            cuda.cuda.synchronize()
            got, got_fft = from_dev_u32(d_out), from_dev_u32(d_fft)
        fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
        want, want_fft, _ = oracle.chain_fixed_words(fft, mag, cfar, apply_window(oracle, p, beats), want_stages=True)
        assert (got_fft == want_fft).all(), win
        assert (got == want).all(), win
    # run-time size 256 of a 1024-point generator: cntWin addresses the first 256 coefficients
    p = chain_params(n, 0, W=16, G=4)
    p.fftParams.windowFunc = WT.Hanning()
    regs8 = dict(regs, fftWin=256)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, 8, 2, regs8)
        d_out = cuda.empty((beats.shape[0], 256), dtype=cuda.int32, device="cuda")
        ch.run(to_dev(cuda, np.ascontiguousarray(beats[:, :256])), beats.shape[0], 1, out=d_out)
        cuda.cuda.synchronize()
        got = from_dev_u32(d_out)
    fft, mag, cfar = oracle_cfgs(oracle, p, 8, 2, regs8)
    want = oracle.chain_fixed_words(fft, mag, cfar, apply_window(oracle, p, np.ascontiguousarray(beats[:, :256])))
    assert (got == want).all()
    # the generic kernels (binPoint 12 takes the 3-multiplier variant, Convergent the fully generic one) and N = 4096 in the fast one
    for bp, trim, log_n in ((12, "RoundHalfUp", 10), (0, "Convergent", 9), (0, "RoundHalfUp", 12)):
        nn = 1 << log_n
        p = chain_params(nn, bp, trim=trim, W=16, G=4)
        p.fftParams.windowFunc = WT.Hamming()
        regs_n = dict(regs, fftWin=nn, thresholdScaler=3 << bp)
        b = np.concatenate([synth.chirps_fixed(4, nn, seed=23), synth.random_beats(2, nn, seed=24)])
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs_n)
            d_out = cuda.empty((b.shape[0], nn), dtype=cuda.int32, device="cuda")
            ch.run(to_dev(cuda, b), b.shape[0], 1, out=d_out)
            cuda.cuda.synchronize()
            got = from_dev_u32(d_out)
        fft, mag, cfar = oracle_cfgs(oracle, p, log_n, 2, regs_n)
        assert (got == oracle.chain_fixed_words(fft, mag, cfar, apply_window(oracle, p, b))).all(), (bp, trim, log_n)
    # IFFT: no window (SDFFFT.scala:236-245)
    p = chain_params(n, 0, W=16, G=4, fftDir=False)
    p.fftParams.windowFunc = WT.Hanning()
    run_and_compare(cuda, oracle, p, 10, 2, regs, beats)


# ------------------------------------------------------------------------------------------------ radix-2 chain (SDFChainRadix2.scala)
@pytest.mark.parametrize("dit", [False, True])
@pytest.mark.parametrize("log_n", [1, 2, 3, 4, 5, 7, 8, 10, 11, 13, 14])
def test_radix2_chain_all_sizes(cuda, oracle, log_n, dit):
    """sdfRadix = "2", DIF and DIT, full-range random beats, every pass geometry of the row-resident kernel; FFT block alone for
    the sizes too small for a CFAR window"""
    n = 1 << log_n
    dec = R.DITDecimType if dit else R.DIFDecimType
    beats = synth.random_beats(23 if log_n < 13 else 4, n, seed=100 + log_n)
    W, G = max(2, min(16, n // 8)), 1
    p = chain_params(n, 0, W=W, G=G, sdfRadix="2", decimType=dec)
    if n > 2 * W + 2 * G + 1:
        regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=max(0, W.bit_length() - 1), peakGrouping=0, cfarMode=dit,
                    windowCells=W, guardCells=G)
        run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)
    fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, {})
    with R.FftMagCfarChainVanilla(p) as ch:
        ch.memWriteWord(p.fftAddress.base, log_n)
        d_in = to_dev(cuda, beats)
        d_f = cuda.empty_like(d_in)
        ch.fft(d_in, beats.shape[0], d_f)
        cuda.cuda.synchronize()
        if n > 2 * W + 2 * G + 1:
            assert ("radix2-dit" if dit else "radix2-dif") in ch.variant
        assert (from_dev_u32(d_f) == oracle.fft_fixed_words(fft, beats)).all()


@pytest.mark.parametrize("dit", [False, True])
@pytest.mark.parametrize("bp,trim,use4,bitrev,fwd,keep", [
    (14, "Convergent", False, True, True, None), (12, "RoundDown", True, False, True, None), (0, "RoundHalfUp", False, False, False, None),
    (12, "RoundHalfToOdd", False, True, True, 0b1011011101), (9, "RoundTowardsZero", False, True, False, 0b0000011111)])
def test_radix2_fft_block_variants(cuda, oracle, dit, bp, trim, use4, bitrev, fwd, keep):
    """radix-2 FFT block alone: binary points, trim types, bit-reversed input (DIT) / output (DIF) without the ping-pong buffer,
    IFFT, keep-LSB stages, run-time sizes (DIF: last k stages, DIT: first k; any k, SDFChainRadix2.scala:127-134)"""
    n, smax = 1024, 10
    kw = dict(sdfRadix="2", decimType=R.DITDecimType if dit else R.DIFDecimType, use4Muls=use4, fftDir=fwd)
    if keep is not None:
        kw["keepMSBorLSB"] = [bool((keep >> i) & 1) for i in range(smax)]
    p = chain_params(n, bp, trim, **kw)
    p.fftParams.useBitReverse = bitrev
    for log_n in (10, 9, 6, 1):
        beats = synth.random_beats(9, 1 << log_n, seed=log_n + 7 * bp, full_range=keep is None)
        fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, {})
        with R.FftMagCfarChainVanilla(p) as ch:
            ch.memWriteWord(p.fftAddress.base, log_n)
            d_in = to_dev(cuda, beats)
            d_f = cuda.empty_like(d_in)
            ch.fft(d_in, beats.shape[0], d_f)
            cuda.cuda.synchronize()
            got = from_dev_u32(d_f)
        want = oracle.fft_fixed_words(fft, beats)
        assert (got == want).all(), f"log_n={log_n}: {(got != want).sum()} beats differ"


@pytest.mark.parametrize("log_n", [1, 2, 3, 4, 6, 7, 9, 10, 11, 12, 14])
def test_radix22_dit_chain_all_sizes(cuda, oracle, log_n):
    """sdfRadix = "2^2" with DITDecimType (SDFChainRadix22.scala:224-352): even and odd stage counts (the twiddle stages swap parity),
    full-range random beats, whole chain and FFT block alone"""
    n = 1 << log_n
    beats = synth.random_beats(21 if log_n < 13 else 4, n, seed=400 + log_n)
    W, G = max(2, min(16, n // 8)), 1
    p = chain_params(n, 0, W=W, G=G, decimType=R.DITDecimType)
    if n > 2 * W + 2 * G + 1:
        regs = dict(fftWin=n, thresholdScaler=3, logOrLinearMode=1, divSum=max(0, W.bit_length() - 1), peakGrouping=1, cfarMode=2,
                    windowCells=W, guardCells=G)
        run_and_compare(cuda, oracle, p, log_n, 2, regs, beats)
    fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, {})
    with R.FftMagCfarChainVanilla(p) as ch:
        ch.memWriteWord(p.fftAddress.base, log_n)
        d_in = to_dev(cuda, beats)
        d_f = cuda.empty_like(d_in)
        ch.fft(d_in, beats.shape[0], d_f)
        cuda.cuda.synchronize()
        if n > 2 * W + 2 * G + 1:
            assert "radix22-dit" in ch.variant
        assert (from_dev_u32(d_f) == oracle.fft_fixed_words(fft, beats)).all()


@pytest.mark.parametrize("numpoints", [1024, 2048])
@pytest.mark.parametrize("bp,trim,use4,bitrev,fwd,keep", [
    (14, "Convergent", False, True, True, None), (12, "RoundDown", True, False, True, None), (0, "RoundHalfUp", False, False, False, None),
    (12, "RoundHalfToOdd", False, True, True, 0b1011011101), (9, "RoundTowardsZero", False, True, False, 0b0000011111)])
def test_radix22_dit_fft_block_variants(cuda, oracle, numpoints, bp, trim, use4, bitrev, fwd, keep):
    """radix-2^2 DIT FFT block: binary points (3-multiplier per-product trim and use4Muls), trim types, bit-reversed input without the
    ping-pong buffer, IFFT, keep-LSB stages, run-time sizes = the FIRST k stages, any k, for an even and an odd elaborated stage count"""
    smax = numpoints.bit_length() - 1
    kw = dict(decimType=R.DITDecimType, use4Muls=use4, fftDir=fwd)
    if keep is not None:
        kw["keepMSBorLSB"] = [bool((keep >> i) & 1) for i in range(smax)]
    p = chain_params(numpoints, bp, trim, **kw)
    p.fftParams.useBitReverse = bitrev
    for log_n in (smax, smax - 1, 6, 5, 1):
        beats = synth.random_beats(7, 1 << log_n, seed=log_n + 11 * bp, full_range=keep is None)
        fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, {})
        with R.FftMagCfarChainVanilla(p) as ch:
            ch.memWriteWord(p.fftAddress.base, log_n)
            d_in = to_dev(cuda, beats)
            d_f = cuda.empty_like(d_in)
            ch.fft(d_in, beats.shape[0], d_f)
            cuda.cuda.synchronize()
            got = from_dev_u32(d_f)
        want = oracle.fft_fixed_words(fft, beats)
        assert (got == want).all(), f"log_n={log_n}: {(got != want).sum()} beats differ"
