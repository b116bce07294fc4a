"""GPU parity, part 2 (through the C ABI): the fp32 flavour of the chain against the fp64 oracle (tolerance 1e-4 relative,
identical detections -- north_star), and the PLFG -> NCO generator kernels against the integer oracle (bit-exact)."""
import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import synth
from rsp_chains_b200.params import (CACFARType, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, GOSCFARType,
                                    MAGParams)
from tests.util import configure, decode_cfar_words, from_dev_u32, oracle_cfgs, to_dev

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4   # north_star: fp32 results within 1e-4 relative error of an fp64 reference


def f32_params(n, W, G, cfar_type=CACFARType):
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True),
        magParams=MAGParams.fixed(),
        cfarParams=CFARParams(leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=cfar_type, sendCut=False))


from oracle.glue import f64_chain, unpack_mask  # noqa: E402  (fp64 dataflow of the DspReal flavour, all three CFAR cores)


@pytest.mark.parametrize("mode", [1, 2, 0])
def test_c3_fp32_chain_go_so(cuda, oracle, mode):
    """BASELINE config 2 (reduced row count for the fp64 oracle): 4 RX x chirps x 2048-point fp32 chain, GO / SO / CA CFAR"""
    n, log_n, W, G, n_rx, n_chirps = 2048, 11, 32, 4, 4, 6
    rows = n_rx * n_chirps
    phases = np.repeat(np.arange(n_rx) * np.pi / 8, n_chirps)
    xc = synth.chirps_complex(rows, n, seed=3, rx_phase=phases)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)      # what the kernel actually sees
    regs = dict(fftWin=n, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)
    scaler = 3.5
    with R.FftMagCfarChainVanilla(f32_params(n, W, G)) as ch:
        configure(ch, log_n, 2, regs)
        ch.set_threshold_scaler_f32(scaler)
        d_in = cuda.from_numpy(x32).cuda()
        d_thr = cuda.empty((rows, n), dtype=cuda.float32, device="cuda")
        d_mask = cuda.zeros((rows, n // 32), dtype=cuda.int32, device="cuda")
        ch.run(d_in, n_chirps, n_rx, out=d_thr, peak_mask=d_mask)
        cuda.cuda.synchronize()
        thr = d_thr.cpu().numpy()
        pk = unpack_mask(d_mask.cpu().numpy().view(np.uint32), n)
        assert ch.variant.startswith("chain_f32")
    mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, scaler, W, G)
    scale = np.abs(want_thr).max()
    rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * scale)
    assert rel.max() < REL_TOL, f"max relative threshold error {rel.max():.3e}"
    # identical detections, except cells whose margin |cut - thr| is itself inside the fp32 tolerance (none in this data)
    margin = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-30)
    decided = margin > 10 * REL_TOL
    assert (pk[decided] == want_pk[decided]).all()
    assert (pk == want_pk).all()
    assert want_pk.sum() >= 2 * rows      # both targets of every row


def test_fp32_blocks_and_sizes(cuda, oracle):
    """fft alone (FFT(x)/N, natural order) for several sizes, magnitude modes, log threshold"""
    rng = np.random.default_rng(1)
    for log_n in (3, 6, 10, 13):
        n = 1 << log_n
        W = max(2, min(16, n // 8)); G = 1
        x = (rng.uniform(-1, 1, (5, n)) + 1j * rng.uniform(-1, 1, (5, n)))
        x32 = np.stack([x.real, x.imag], axis=-1).astype(np.float32)
        with R.FftMagCfarChainVanilla(f32_params(n, W, G)) as ch:
            configure(ch, log_n, 2, dict(fftWin=n, logOrLinearMode=0, divSum=int(np.log2(W)), cfarMode=0, windowCells=W, guardCells=G))
            d_in = cuda.from_numpy(x32).cuda()
            d_out = cuda.empty((5, n, 2), dtype=cuda.float32, device="cuda")
            ch.fft(d_in, 5, d_out)
            cuda.cuda.synchronize()
            got = d_out.cpu().numpy()
        xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
        want = np.fft.fft(xq, axis=1) / n
        err = np.abs(got[..., 0] + 1j * got[..., 1] - want).max() / np.abs(want).max()
        assert err < REL_TOL, f"N={n}: {err:.2e}"


@pytest.mark.parametrize("log_n,W,G", [(10, 16, 2), (10, 64, 8), (11, 32, 4), (12, 64, 4), (13, 16, 1), (14, 64, 8)])
def test_fast_f32_kernel_variants(cuda, oracle, log_n, W, G):
    """chain_f32_fast<S>: every size it is built for, CA / GO / SO x linear / log threshold, ragged row counts (two chirps per CTA at
    N = 1024, edge iterations with W + G > 64), rows 60 dB apart in level, thresholds + peak mask + detection list"""
    n = 1 << log_n
    rows = 5 if log_n <= 12 else 3
    xc = synth.chirps_complex(rows, n, seed=10 + log_n)
    xc[1, :] *= 1e-3                                   # a weak row
    xc[2, :] *= 1.9                                    # close to full scale
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    for mode, lin, scaler in ((0, 1, 3.5), (1, 1, 2.5), (2, 1, 4.0), (0, 0, 1e-4), (2, 0, 0.0)):
        regs = dict(fftWin=n, logOrLinearMode=lin, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)
        with R.FftMagCfarChainVanilla(f32_params(n, max(W, 16), max(G, 1))) as ch:
            configure(ch, log_n, 2, regs)
            ch.set_threshold_scaler_f32(scaler)
            assert ch.variant == f"chain_f32_fast<{log_n}>", ch.variant
            d_in = cuda.from_numpy(x32).cuda()
            d_thr = cuda.empty((rows, n), dtype=cuda.float32, device="cuda")
            d_mask = cuda.zeros((rows, n // 32), dtype=cuda.int32, device="cuda")
            det_list = cuda.zeros((rows * n, 4), dtype=cuda.int32, device="cuda")
            det_count = cuda.zeros(1, dtype=cuda.int32, device="cuda")
            ch.run(d_in, rows, 1, out=d_thr, peak_mask=d_mask, det_list=det_list, det_count=det_count)
            cuda.cuda.synchronize()
            thr = d_thr.cpu().numpy()
            pk = unpack_mask(d_mask.cpu().numpy().view(np.uint32), n)
            k = int(det_count.item())
            dets = det_list[:k].cpu().numpy()
        mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, scaler, W, G)
        scale = np.abs(want_thr).max(axis=1, keepdims=True)
        rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-6 * scale)
        assert rel.max() < REL_TOL, f"mode {mode} lin {lin}: max relative threshold error {rel.max():.3e}"
        margin = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-30)
        decided = margin > 10 * REL_TOL
        assert (pk[decided] == want_pk[decided]).all(), f"mode {mode} lin {lin}"
        # the detection list is the peak mask, with the threshold and the cell under test as float bits
        assert k == int(pk.sum())
        got = set(map(tuple, dets[:, :2].tolist()))
        assert got == set(map(tuple, np.argwhere(pk == 1).tolist()))
        if k:
            r0, b0 = dets[0, 0], dets[0, 1]
            assert dets[0, 2:].view(np.float32)[0] == thr[r0, b0]


def test_fp32_window_functions(cuda, oracle):
    """windowFunc in the fp32 flavour (protoWin = DspReal, SDFFFTParams.scala:177): x * window(n) in front of the forward transform"""
    from rsp_chains_b200 import WindowFunctionTypes as WT
    n, log_n, W, G, rows = 2048, 11, 32, 4, 4
    xc = synth.chirps_complex(rows, n, seed=31)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    regs = dict(fftWin=n, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    for win in (WT.Hanning(), WT.Blackman()):
        p = f32_params(n, W, G)
        p.fftParams.windowFunc = win
        w = oracle.window_coeffs(win.kind, n, win.a0, win.a1, win.a2)
        xq = (x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)) * w.astype(np.float32).astype(np.float64)[None, :]
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs)
            ch.set_threshold_scaler_f32(3.0)
            assert ch.variant == f"chain_f32_fast<{log_n}>", ch.variant
            d_thr = cuda.empty((rows, n), dtype=cuda.float32, device="cuda")
            d_mask = cuda.zeros((rows, n // 32), dtype=cuda.int32, device="cuda")
            ch.run(cuda.from_numpy(x32).cuda(), rows, 1, out=d_thr, peak_mask=d_mask)
            cuda.cuda.synchronize()
            thr = d_thr.cpu().numpy()
            pk = unpack_mask(d_mask.cpu().numpy().view(np.uint32), n)
        mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, 3.0, W, G)
        rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * np.abs(want_thr).max())
        assert rel.max() < REL_TOL, f"{win.kind}: {rel.max():.3e}"
        decided = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-30) > 10 * REL_TOL
        assert (pk[decided] == want_pk[decided]).all()


# ---------------------------------------------------------------------------------------------------- PLFG -> NCO

def write_plfg(ch, p, oracle, starting_point, segs, n_rep=1, frame_num=1, inter=10):
    """program the PLFG through the register bus as the reference tester does (RspChainVanillaTester.scala:86-130)"""
    bb = p.beatBytes
    for i, (slope, ns, up, rst) in enumerate(segs):
        ch.memWriteWord(p.plfgRAM.base + i * bb, oracle.plfg_ram_word(slope, ns, up, rst))
    seg0 = 6
    rep0 = seg0 + p.plfgParams.maxChirpOrdinalNum
    ord0 = rep0 + p.plfgParams.maxNumOfDifferentChirps
    ch.memWriteWord(p.plfgAddress.base + 2 * bb, frame_num)
    ch.memWriteWord(p.plfgAddress.base + 3 * bb, inter)
    ch.memWriteWord(p.plfgAddress.base + 4 * bb, 1)
    ch.memWriteWord(p.plfgAddress.base + 5 * bb, starting_point)
    ch.memWriteWord(p.plfgAddress.base + seg0 * bb, len(segs))
    ch.memWriteWord(p.plfgAddress.base + rep0 * bb, n_rep)
    ch.memWriteWord(p.plfgAddress.base + ord0 * bb, 0)
    ch.memWriteWord(p.plfgAddress.base + 1 * bb, 0)
    ch.memWriteWord(p.plfgAddress.base + 0 * bb, 1)   # enable


def test_generator_matches_oracle_and_c1_chain(cuda, oracle):
    """BASELINE config 0, reference-faithful reading: PLFG (1 segment, slope 0, startingPoint 16) -> NCO (table 128, 16 bit,
    phase 9) -> 1024-point fixed FFT -> JPL -> CA-CFAR: tone at bin 16 * 1024 / (4 * 128) = 32
    (src/test/scala/RspChainVanillaTester.scala:85-94).  Then a triangle chirp (up / down segments) for the closed-form phase."""
    p = R.rsp_chain_vanilla_app_params()
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    # segmentNums is log2Ceil(maxNumOfSegments) = 2 bits wide (PLFGDspBlockMem.scala:49,161): at most 3 segments per chirp here
    for name, start, segs, n_rep in (("tone", 16, [(0, 255, True, False)] * 3, 1),
                                     ("triangle", 3, [(1, 200, True, False), (2, 180, False, False), (0, 100, False, True)], 2)):
        n_per_chirp = sum(s[1] for s in segs)
        with R.RspChainVanilla(p) as ch:
            configure(ch, 10, 2, regs)
            write_plfg(ch, p, oracle, start, segs, n_rep=n_rep, frame_num=0, inter=7)
            n_samples = 4 * 1024
            d_nco = cuda.zeros(n_samples, dtype=cuda.int32, device="cuda")
            d_plfg = cuda.zeros(n_samples, dtype=cuda.int32, device="cuda")
            ch.generate(n_samples, d_nco, d_plfg)
            d_out = cuda.zeros((4, 1024), dtype=cuda.int32, device="cuda")
            ch.run_from_generator(4, 1, out=d_out)
            cuda.cuda.synchronize()
            got_nco, got_plfg, got_out = from_dev_u32(d_nco), d_plfg.cpu().numpy(), from_dev_u32(d_out)
        ram = [oracle.plfg_ram_word(*s) for s in segs]
        op = oracle.Plfg(frame_num=0, interframe_num_of_samples=7, different_chirps_num=1, starting_point=start,
                         segment_nums=[len(segs)], repeated_chirp_nums=[n_rep], chirp_ordinal_nums=[0], ram=ram)
        want_plfg = op.run(n_samples)
        assert (got_plfg == want_plfg).all(), name
        want_nco, _ = oracle.nco_words(want_plfg, 128, 16, 9)
        assert (got_nco == want_nco).all(), name
        fft, mag, cfar = oracle_cfgs(oracle, p, 10, 2, regs)
        want_out = oracle.chain_fixed_words(fft, mag, cfar, want_nco.reshape(4, 1024))
        assert (got_out == want_out).all(), name
        if name == "tone":
            _, bins, peak = decode_cfar_words(got_out, 1024)
            assert n_per_chirp == 3 * 255
            assert peak[0, 32] == 1     # (quantisation-noise bins next to an all-zero window may flag too: threshold 0)


def test_fp32_chain_accepts_radix2(cuda, oracle):
    """sdfRadix = "2" (DIF) in the DspReal flavour: the same FFT(x)/N within 1e-4 of the fp64 dataflow, identical detections, also at a
    run-time size that drops an odd number of stages (legal for the radix-2 chain only, SDFChainRadix2.scala:127-134); DIT -> unsupported"""
    n, W, G = 1024, 16, 2
    p = f32_params(n, W, G)
    p.fftParams.sdfRadix = "2"
    for log_n in (10, 9):
        m = 1 << log_n
        xc = synth.chirps_complex(5, m, seed=log_n)
        x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
        xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
        regs = dict(fftWin=m, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
        with R.FftMagCfarChainVanilla(p) as ch:
            configure(ch, log_n, 2, regs)
            ch.set_threshold_scaler_f32(3.5)
            d_thr = cuda.empty((5, m), dtype=cuda.float32, device="cuda")
            d_mask = cuda.zeros((5, max(1, m // 32)), dtype=cuda.int32, device="cuda")
            ch.run(cuda.from_numpy(x32).cuda(), 5, 1, out=d_thr, peak_mask=d_mask)
            cuda.cuda.synchronize()
            thr = d_thr.cpu().numpy()
            pk = unpack_mask(d_mask.cpu().numpy().view(np.uint32), m)
        mag, want_thr, want_pk = f64_chain(oracle, xq, log_n, 2, regs, 3.5, W, G)
        rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * np.abs(want_thr).max())
        assert rel.max() < REL_TOL
        sure = np.abs(mag - want_thr) / np.maximum(np.abs(want_thr), 1e-30) > 10 * REL_TOL
        assert (pk[sure] == want_pk[sure]).all()


@pytest.mark.parametrize("radix", ["2", "2^2"])
def test_fp32_dit_orders(cuda, oracle, radix):
    """DITDecimType in the DspReal flavour: same FFT(x)/N; with useBitReverse natural in / natural out (fast kernel), without it the
    beats arrive bit-reversed and the bins leave in natural order (SDFFFT.scala:217-270)"""
    n, log_n = 2048, 11
    rev = np.array([int(format(i, f"0{log_n}b")[::-1], 2) for i in range(n)])
    xc = synth.chirps_complex(4, n, seed=21)
    x32 = np.stack([xc.real, xc.imag], axis=-1).astype(np.float32)
    want = np.fft.fft(x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64), axis=1) / n
    for use_br in (True, False):
        p = f32_params(n, 32, 4)
        p.fftParams.sdfRadix, p.fftParams.decimType, p.fftParams.useBitReverse = radix, R.DITDecimType, use_br
        with R.FftMagCfarChainVanilla(p) as ch:
            ch.memWriteWord(p.fftAddress.base, log_n)
            d_in = cuda.from_numpy(np.ascontiguousarray(x32 if use_br else x32[:, rev])).cuda()
            d_out = cuda.empty_like(d_in)
            ch.fft(d_in, 4, d_out)
            cuda.cuda.synchronize()
            got = d_out.cpu().numpy().astype(np.float64)
        err = np.abs((got[..., 0] + 1j * got[..., 1]) - want).max() / np.abs(want).max()
        assert err < REL_TOL, f"useBitReverse={use_br}: {err:.2e}"
