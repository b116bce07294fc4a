"""Slow-time dimension (not in the reference): the range-Doppler chain built from the reference's blocks against the same
composition of the oracle's blocks -- oracle FFT along fast time, numpy transpose, oracle FFT along slow time, oracle LogMagMux per RX,
integer mean, oracle CFAR -- bit for bit; rsp_corner_turn on ragged shapes and every element size; rsp_mag_combine_run argument checks."""
import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import _abi, synth
from tests.test_gpu_chain import chain_params
from tests.util import configure, from_dev_u32, oracle_cfgs, to_dev

pytestmark = pytest.mark.gpu


def _sx16(v):
    return ((v.astype(np.int64) & 0xFFFF) ^ 0x8000) - 0x8000


def _frame(n_rx, m, n, seed):
    """n_rx x m chirps of n samples: two targets whose phase advances from chirp to chirp (Doppler bins m/8 and 3m/4) + noise"""
    rng = np.random.default_rng(seed)
    t = np.arange(n)[None, None, :] / n
    k = np.arange(m)[None, :, None]
    rx = np.arange(n_rx)[:, None, None]
    x = 0.35 * np.exp(2j * np.pi * (0.125 * n * t + k / 8 + rx / 16)) + 0.2 * np.exp(2j * np.pi * (0.25 * n * t + 3 * k / 4 - rx / 8))
    x = x + (rng.normal(size=x.shape) + 1j * rng.normal(size=x.shape)) * 2.0 ** -6
    re = np.clip(np.round(x.real * 8192), -32768, 32767).astype(np.int64)
    im = np.clip(np.round(x.imag * 8192), -32768, 32767).astype(np.int64)
    return (((re & 0xFFFF) << 16) | (im & 0xFFFF)).astype(np.uint32)


@pytest.mark.parametrize("cfar_axis", ["doppler", "range"])
@pytest.mark.parametrize("n_rx,m,n,sel", [(4, 64, 256, 2), (2, 128, 1024, 2), (1, 32, 64, 1), (8, 256, 128, 3)])
def test_range_doppler_chain_equals_the_composed_oracle(cuda, oracle, n_rx, m, n, sel, cfar_axis):
    W, G = (8, 2) if min(n, m) <= 64 else (16, 2)
    pr, pd = chain_params(n, 0, W=W, G=G), chain_params(m, 0, W=W, G=G)
    nc = m if cfar_axis == "doppler" else n
    regs = dict(fftWin=nc, thresholdScaler=5, logOrLinearMode=1, divSum=int(np.log2(W)), peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    got = _check_frame(cuda, oracle, pr, pd, n_rx, m, n, sel, cfar_axis, regs, _frame(n_rx, m, n, seed=n + m))
    if sel == 2:     # both targets are found: range bins n/8 and n/4, Doppler bins m/8 and 3m/4
        peak = (got & 1).astype(bool)
        peak = peak if cfar_axis == "doppler" else peak.T
        assert peak[n // 8, m // 8] and peak[n // 4, (3 * m) // 4]


def test_randomised_range_doppler_frames(cuda, oracle):
    """16 seeded random frames: 1 ... 8 RX, 8 ... 512 chirps, 64 ... 2048 samples, binPoint 0 / 12, CA / GOS / combined-window OS along either axis,
    the three magnitudes, linear / log -- each stage and the result equal to the composed oracle blocks"""
    import os
    from rsp_chains_b200.params import CACFARType, GOSCFARType, OSGoldenCFARType
    rng = np.random.default_rng(int(os.environ.get("RSP_FUZZ_SEED", "555")))
    for trial in range(int(os.environ.get("RSP_FUZZ_TRIALS", "16"))):
        n_rx = 1 << int(rng.integers(0, 4))
        m, n = 1 << int(rng.integers(3, 10)), 1 << int(rng.integers(6, 12))
        cfar_axis = "doppler" if rng.random() < 0.5 else "range"
        nc = m if cfar_axis == "doppler" else n
        bp = int(rng.choice([0, 12]))
        ctype = [CACFARType, GOSCFARType, OSGoldenCFARType][int(rng.integers(0, 3))]
        W = 1 << int(rng.integers(1, max(2, min(5, int(np.log2(nc)) - 2))))
        G = int(rng.integers(1, 3))
        if nc <= 2 * W + 2 * G + 1:
            continue
        sel = int(rng.integers(1, 4))      # (0 is the bypass: nothing to combine, refused)
        lin = int(rng.integers(0, 2)) if sel != 3 else 0
        pr, pd = chain_params(n, bp, cfar_type=ctype, W=max(W, 4), G=G), chain_params(m, bp, cfar_type=ctype, W=max(W, 4), G=G)
        regs = dict(fftWin=nc, thresholdScaler=int(4 * (1 << bp)) if lin else int(rng.integers(0, 30)), logOrLinearMode=lin, divSum=int(np.log2(W)),
                    peakGrouping=0, cfarMode=int(rng.integers(0, 3)), windowCells=W, guardCells=G, indexLagg=int(rng.integers(1, W + 1)),
                    indexLead=int(rng.integers(1, W + 1)))
        _check_frame(cuda, oracle, pr, pd, n_rx, m, n, sel, cfar_axis, regs, _frame(n_rx, m, n, seed=700 + trial),
                     what=(trial, dict(n_rx=n_rx, m=m, n=n, axis=cfar_axis, bp=bp, ctype=ctype, W=W, G=G, sel=sel, lin=lin)))


def _check_frame(cuda, oracle, pr, pd, n_rx, m, n, sel, cfar_axis, regs, frame, what=None):
    log_n, log_m = n.bit_length() - 1, m.bit_length() - 1
    # the oracle's blocks, composed
    fr, _, cr = oracle_cfgs(oracle, pr, log_n, sel, regs if cfar_axis == "range" else dict(regs, fftWin=n))
    fd, md, cd = oracle_cfgs(oracle, pd, log_m, sel, regs if cfar_axis == "doppler" else dict(regs, fftWin=m))
    want_range = oracle.fft_fixed_words(fr, frame.reshape(n_rx * m, n)).reshape(n_rx, m, n)
    turned = np.ascontiguousarray(want_range.transpose(0, 2, 1))
    want_dopp = oracle.fft_fixed_words(fd, turned.reshape(n_rx * n, m)).reshape(n_rx, n, m)
    mags = oracle.mag_words(md, want_dopp.reshape(-1)).reshape(n_rx, n, m)
    mean = _sx16(mags).sum(axis=0) >> (n_rx.bit_length() - 1)
    want_comb = (mean & 0xFFFF).astype(np.uint32)
    want = oracle.cfar_words(cd, want_comb) if cfar_axis == "doppler" else oracle.cfar_words(cr, np.ascontiguousarray(want_comb.T))
    with R.RangeDopplerChain(pr, pd, n_rx, cfar_axis=cfar_axis) as rd:
        configure(rd.range, log_n, sel, regs if cfar_axis == "range" else dict(regs, fftWin=n))
        configure(rd.doppler, log_m, sel, regs if cfar_axis == "doppler" else dict(regs, fftWin=m))
        d_in = to_dev(cuda, frame.reshape(n_rx * m, n)).view(n_rx, m, n)
        st = rd.stages(d_in)
        cuda.cuda.synchronize()
        assert (from_dev_u32(st["range_fft"]).reshape(n_rx, m, n) == want_range).all(), what
        assert (from_dev_u32(st["turned"]).reshape(n_rx, n, m) == turned).all(), what
        assert (from_dev_u32(st["doppler_fft"]).reshape(n_rx, n, m) == want_dopp).all(), what
        assert (from_dev_u32(st["combined"]).reshape(n, m) == want_comb).all(), what
        rows, cols = (n, m) if cfar_axis == "doppler" else (m, n)
        d_out = cuda.empty((rows, cols), dtype=cuda.int32, device="cuda")
        launches = rd.kernel_launches
        rd.run(d_in, out=d_out)
        cuda.cuda.synchronize()
        assert rd.kernel_launches - launches == (5 if cfar_axis == "doppler" else 6)
        got = from_dev_u32(d_out)
    assert (got == want).all(), what
    return got


@pytest.mark.parametrize("dtype,rows,cols,batch", [("int16", 100, 37, 3), ("int32", 64, 64, 1), ("int32", 33, 1000, 2), ("int64", 17, 130, 5),
                                                   ("int16", 4096, 256, 1), ("int32", 1, 77, 4), ("int32", 500, 1, 2)])
def test_corner_turn_shapes_and_element_sizes(cuda, dtype, rows, cols, batch):
    p = chain_params(64, 0, W=8, G=2)
    rng = np.random.default_rng(rows * cols)
    x = rng.integers(-2 ** 15, 2 ** 15, (batch, rows, cols)).astype(dtype)
    with R.FftMagCfarChainVanilla(p) as ch:
        d_in = cuda.from_numpy(x).cuda()
        d_out = cuda.empty((batch, cols, rows), dtype=d_in.dtype, device="cuda")
        ch.corner_turn(d_in, rows, cols, d_out, batch=batch)
        cuda.cuda.synchronize()
        assert (d_out.cpu().numpy() == x.transpose(0, 2, 1)).all()
        # twice is the identity
        d_back = cuda.empty_like(d_in)
        ch.corner_turn(d_out, cols, rows, d_back, batch=batch)
        cuda.cuda.synchronize()
        assert bool((d_back == d_in).all().item())
        st = _abi.lib().rsp_corner_turn(ch._h, d_in.data_ptr(), d_in.data_ptr(), rows, cols, batch, x.itemsize, None)
        assert st == _abi.RSP_EINVAL
        st = _abi.lib().rsp_corner_turn(ch._h, d_in.data_ptr(), d_out.data_ptr(), rows, cols, batch, 3, None)
        assert st == _abi.RSP_EINVAL


def test_mag_combine_fp32_and_argument_checks(cuda, oracle):
    from tests.test_gpu_f32_gen import f32_params
    n, n_rx = 256, 4
    rng = np.random.default_rng(9)
    x = rng.uniform(-1, 1, (n_rx, 5 * n, 2)).astype(np.float32)
    with R.FftMagCfarChainVanilla(f32_params(n, 16, 2)) as ch:
        configure(ch, 8, 2, dict(fftWin=n, logOrLinearMode=1, divSum=4, cfarMode=0, windowCells=16, guardCells=2))
        d_in = cuda.from_numpy(x).cuda()
        d_out = cuda.empty(5 * n, dtype=cuda.float32, device="cuda")
        ch.mag_combine(d_in, 5 * n, n_rx, d_out)
        cuda.cuda.synchronize()
        re, im = np.abs(x[..., 0].astype(np.float64)), np.abs(x[..., 1].astype(np.float64))
        u, v = np.maximum(re, im), np.minimum(re, im)
        want = np.maximum(u + v / 8, 7 * u / 8 + v / 2).mean(axis=0)
        assert np.abs(d_out.cpu().numpy() - want).max() < 1e-5
        assert _abi.lib().rsp_mag_combine_run(ch._h, d_in.data_ptr(), 5 * n, 3, d_out.data_ptr(), None) == _abi.RSP_EINVAL
        ch.memWriteWord(ch.params.magAddress.base, 0)
        assert _abi.lib().rsp_mag_combine_run(ch._h, d_in.data_ptr(), 5 * n, 4, d_out.data_ptr(), None) == _abi.RSP_EINVAL
        assert "bypass" in _abi.lib().rsp_last_error().decode()


@pytest.mark.parametrize("log_n", [10, 11, 12, 13, 14])
def test_fft_block_alone_on_the_fast_passes(cuda, oracle, log_n):
    """rsp_fft_run in the fast chain's FFT configuration runs fft_fixed_fast_kernel<S> (the fast passes + a bit-reversed scatter of the
    output beats): quiet rows, full-scale random rows (every wrap, the -32768 corners), a DC row at full scale, ragged row counts
    (two chirps per CTA at N = 1024), a window; bit for bit against the oracle and against the run-time-parameter kernel."""
    import os
    n = 1 << log_n
    rows = 13 if log_n <= 12 else 5
    beats = np.concatenate([synth.chirps_fixed(rows - 4, n, seed=40 + log_n), synth.random_beats(3, n, seed=60 + log_n),
                            np.full((1, n), 0x7FFF8000, dtype=np.uint32)])
    p = chain_params(n, 0, W=16, G=2)
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=2)
    fft, _, _ = oracle_cfgs(oracle, p, log_n, 2, regs)
    want = oracle.fft_fixed_words(fft, beats)
    with R.FftMagCfarChainVanilla(p) as ch:
        configure(ch, log_n, 2, regs)
        d_in = to_dev(cuda, beats)
        d_out = cuda.zeros_like(d_in)
        ch.fft(d_in, beats.shape[0], d_out)
        cuda.cuda.synchronize()
        assert (from_dev_u32(d_out) == want).all()
        for k in (1, 2, 3):      # ragged counts
            d_out.zero_()
            ch.fft(d_in, k, d_out)
            cuda.cuda.synchronize()
            got = from_dev_u32(d_out)
            assert (got[:k] == want[:k]).all() and not got[k:].any()
    # with a window (the product multiplies, the oracle gets the windowed beats)
    from oracle.glue import apply_window
    from rsp_chains_b200.params import WindowFunctionTypes as WT
    pw = chain_params(n, 0, W=16, G=2)
    pw.fftParams.windowFunc = WT.Hamming(16, 0.54, 0.46)
    want_w = oracle.fft_fixed_words(fft, apply_window(oracle, pw, beats))
    with R.FftMagCfarChainVanilla(pw) as ch:
        configure(ch, log_n, 2, regs)
        d_out = cuda.zeros_like(d_in)
        ch.fft(d_in, beats.shape[0], d_out)
        cuda.cuda.synchronize()
        assert (from_dev_u32(d_out) == want_w).all()


def test_range_doppler_chain_fp32_flavour(cuda, oracle):
    """the same chain in the DspReal flavour: float2 beats through both transforms (FFT(x)/N each), the 8-byte corner turn, JPL magnitude
    + RX mean in floats, CA-CFAR along Doppler; 1e-4 relative against the fp64 dataflow and identical decided detections"""
    from tests.test_gpu_f32_gen import f32_params
    from oracle.glue import unpack_mask
    n_rx, m, n, W, G, scaler = 4, 64, 256, 16, 2, 4.0
    rng = np.random.default_rng(3)
    t = np.arange(n)[None, None, :] / n
    k = np.arange(m)[None, :, None]
    rx = np.arange(n_rx)[:, None, None]
    x = 0.35 * np.exp(2j * np.pi * (32 * t + k / 8 + rx / 16)) + 0.2 * np.exp(2j * np.pi * (64 * t + 3 * k / 4 - rx / 8))
    x = x + (rng.normal(size=x.shape) + 1j * rng.normal(size=x.shape)) * 2.0 ** -6
    x32 = np.stack([x.real, x.imag], axis=-1).astype(np.float32)
    xq = x32[..., 0].astype(np.float64) + 1j * x32[..., 1].astype(np.float64)
    regs = dict(fftWin=m, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    with R.RangeDopplerChain(f32_params(n, W, G), f32_params(m, W, G), n_rx) as rd:
        configure(rd.range, 8, 2, dict(regs, fftWin=n))
        configure(rd.doppler, 6, 2, regs)
        rd.doppler.set_threshold_scaler_f32(scaler)
        d_in = cuda.from_numpy(x32).cuda()
        d_thr = cuda.zeros((n, m), dtype=cuda.float32, device="cuda")
        d_mask = cuda.zeros((n, m // 32), dtype=cuda.int32, device="cuda")
        st = rd.stages(d_in)
        rd.run(d_in, out=d_thr, peak_mask=d_mask)
        cuda.cuda.synchronize()
        comb = st["combined"].cpu().numpy().reshape(n, m)
        thr = d_thr.cpu().numpy()
        pk = unpack_mask(d_mask.cpu().numpy().view(np.uint32), m)
    rd64 = np.fft.fft(np.fft.fft(xq, axis=2) / n, axis=1) / m                      # [rx][doppler][range]
    re, im = np.abs(rd64.real), np.abs(rd64.imag)
    u, v = np.maximum(re, im), np.minimum(re, im)
    want_comb = np.maximum(u + v / 8, 7 * u / 8 + v / 2).mean(axis=0).T            # [range][doppler]
    scale = want_comb.max()
    assert (np.abs(comb - want_comb) / np.maximum(want_comb, 1e-3 * scale)).max() < 1e-4
    cfg = oracle.cfar_cfg(m, W, G, 0, core=oracle.CORE_MEM, cfar_mode=0, div_sum=4, log_or_linear=1, peak_grouping=0, guard_window_size=G)
    want_thr = np.zeros((n, m)); want_pk = np.zeros((n, m), np.uint8)
    for r in range(n):
        want_thr[r], want_pk[r] = oracle.cfar_f64(cfg, scaler, want_comb[r])
    rel = np.abs(thr - want_thr) / np.maximum(np.abs(want_thr), 1e-3 * np.abs(want_thr).max())
    assert rel.max() < 1e-4, rel.max()
    decided = np.abs(want_comb - want_thr) / np.maximum(np.abs(want_thr), 1e-30) > 1e-3
    assert (pk[decided] == want_pk[decided]).all()
    assert pk[32, m // 8] and pk[64, (3 * m) // 4]
