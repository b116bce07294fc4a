"""Pins the oracle's fixed-point FFT and LogMagMux restatement with every criterion the reference's own tests apply:
 - FixedPointSDFFFTTester.simpleTestFFT (generators/sdf-fft/src/test/scala/SDFFFTTester.scala:27-71): every output bin within
   fixTolLSBs t of fourierTr(x) / 2^(#scaled stages), i.e. |raw error| <= 2^t - 1, on genRandSignal inputs (U[0,1) re/im,
   TesterUtil.scala:141-147), over the Radix22RandomSpec grid (Radix22RandomSpec.scala:17-44: N = 2..1024, 16-bit Q2.14,
   Convergent, scale every stage, t = 3)
 - the published SQNR curve (generators/sdf-fft/doc/images/SQNRdataWidth.png, producer Radix22SQNRSpec.scala:40-173, metric
   TesterUtil.scala:44-54)
 - checkFFTError |delta| <= 4 (TesterUtil.scala:222-228) on the AXI-level tone test
 - LogMagMuxSpec closed forms (generators/logMagMux/src/test/scala/LogMagMuxSpec.scala:21-31), t = 3 (JPL) / 7 (|z|^2)
The bit-level results themselves are unpinned by the reference (it ships no vectors); see DESIGN.md."""
import numpy as np
import pytest


def java_rand_signal(oracle, n, seed):
    """genRandSignal: Complex(Random.nextDouble(), Random.nextDouble()) per sample, scala.util.Random = java.util.Random"""
    r = oracle.JavaRandom(seed)
    re, im = np.zeros(n), np.zeros(n)
    for i in range(n):
        re[i] = r.next_double()
        im[i] = r.next_double()
    return re, im


def quantize(x, bp):   # dsptools poke of a Double into a FixedPoint port: round(x * 2^bp)
    return np.floor(x * (1 << bp) + 0.5).astype(np.int64)


@pytest.mark.parametrize("n", [2, 4, 8, 16, 32, 64, 128, 256, 512, 1024])
@pytest.mark.parametrize("trim,tol", [("Convergent", 3), ("RoundHalfUp", 3)])
def test_fft_radix22_random_spec_tolerance(oracle, n, trim, tol):
    log_n = n.bit_length() - 1
    cfg = oracle.fft_cfg(log_n, 16, 14, 16, trim, False, True, (1 << log_n) - 1, True)
    for rep in range(2):                                  # numSamples = 2 (Radix22RandomSpec.scala:16)
        re, im = java_rand_signal(oracle, n, 11110 + rep)
        qr, qi = quantize(re, 14), quantize(im, 14)
        ore, oim = oracle.fft_fixed(cfg, qr, qi)
        want = np.fft.fft(re + 1j * im) / n * (1 << 14)    # out / scalingFactor, in raw LSB
        err = max(np.abs(ore - want.real).max(), np.abs(oim - want.imag).max())
        assert err <= (1 << tol) - 1, f"N={n} {trim}: max raw error {err:.2f} LSB"


def sqnr_db(got, want):   # calc_sqnr, TesterUtil.scala:44-54 (sums of magnitudes, 20 log10)
    return 20 * np.log10(np.abs(want).sum() / np.abs(got - want).sum())


def test_fft_reproduces_published_sqnr_curve(oracle):
    """16-bit, scale every stage: ~84 dB at N=2 falling ~linearly to ~50 dB at N=1024 (Convergent),
    ~46.5 dB at N=1024 for RoundHalfUp (SQNRroundings.png).  Median over >= 8 random signals, +-2.5 dB."""
    published = {2: 84, 4: 80, 8: 75, 16: 71, 32: 67, 64: 63.5, 128: 60, 256: 56.5, 512: 53, 1024: 50}
    for n, want_db in published.items():
        log_n = n.bit_length() - 1
        cfg = oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", False, True, (1 << log_n) - 1, True)
        vals = []
        for rep in range(max(8, 512 // n)):                 # tiny transforms have few error samples: average more of them
            re, im = java_rand_signal(oracle, n, 100 + rep)
            ore, oim = oracle.fft_fixed(cfg, quantize(re, 14), quantize(im, 14))
            ref = np.fft.fft(re + 1j * im) / n
            vals.append(sqnr_db((ore + 1j * oim) / (1 << 14), ref))
        assert abs(np.median(vals) - want_db) < 2.5, f"N={n}: {np.median(vals):.1f} dB vs published ~{want_db}"
    cfg = oracle.fft_cfg(10, 16, 14, 16, "RoundHalfUp", False, True, 1023, True)
    vals = []
    for rep in range(8):
        re, im = java_rand_signal(oracle, 1024, 100 + rep)
        ore, oim = oracle.fft_fixed(cfg, quantize(re, 14), quantize(im, 14))
        vals.append(sqnr_db((ore + 1j * oim) / (1 << 14), np.fft.fft(re + 1j * im) / 1024))
    assert abs(np.mean(vals) - 46.5) < 2.5


@pytest.mark.parametrize("n", [64, 1024, 2048, 16384])
def test_fft_bp0_tone_axi_level(oracle, n):
    """integer beats (binPoint 0) as the chain apps use (src/main/scala/RspChain.scala:118): checkFFTError |delta| <= 4
    against fft / N (TesterUtil.scala:222-228, FFTBlockWithWindowingSpec.scala:149)"""
    log_n = n.bit_length() - 1
    t = np.arange(n)
    x = 0.4 * np.exp(2j * np.pi * (n // 8) * t / n) + 0.2 * np.exp(2j * np.pi * (n // 4) * t / n)
    qr, qi = np.floor(x.real * 8192).astype(np.int64), np.floor(x.imag * 8192).astype(np.int64)
    cfg = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, (1 << log_n) - 1, True)
    ore, oim = oracle.fft_fixed(cfg, qr, qi)
    want = np.fft.fft(qr + 1j * qi) / n
    assert np.abs(ore - want.real).max() <= 4 and np.abs(oim - want.imag).max() <= 4
    assert int(np.abs(ore + 1j * oim).argmax()) == n // 8
    # packed-beat entry point agrees with the array entry point
    words = oracle.fft_fixed_words(cfg, oracle.pack_iq(qr, qi))
    r2, i2 = oracle.unpack_iq(words)
    assert (r2 == ore).all() and (i2 == oim).all()


def test_fft_fixed_matches_f64_flavour(oracle):
    """the DspReal flavour of the same dataflow equals numpy fft / N to 1e-12; the fixed one stays within the tolerance of it"""
    rng = np.random.default_rng(3)
    for log_n in (1, 2, 3, 5, 8, 11):
        n = 1 << log_n
        re, im = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
        fr, fi = oracle.fft_f64(log_n, re, im)
        ref = np.fft.fft(re + 1j * im) / n
        assert np.abs(fr + 1j * fi - ref).max() < 1e-12
        br, bi = oracle.fft_f64(log_n, re, im, bit_reverse=False)     # DIF order = bit-reversed bins
        rev = np.array([int(format(i, f"0{log_n}b")[::-1], 2) for i in range(n)])
        assert np.abs(br + 1j * bi - ref[rev]).max() < 1e-12


def test_fft_runtime_size_is_last_k_stages(oracle):
    """run-time size N' = 2^k equals a k-stage transform (SDFChainRadix22.scala:146-150,216-219): keep_msb_mask bits are those
    of the LAST k physical stages, which the glue shifts down"""
    rng = np.random.default_rng(5)
    re, im = rng.integers(-8000, 8000, 256), rng.integers(-8000, 8000, 256)
    full = oracle.fft_cfg(8, 16, 0, 16, "RoundHalfUp", False, True, 0xFF, True)
    a = oracle.fft_fixed(full, re, im)
    want = np.fft.fft(re + 1j * im) / 256
    assert np.abs(a[0] - want.real).max() <= 4


def test_fft_keep_lsb_and_ifft(oracle):
    """keepMSBorLSB(i) = false keeps the LSBs (no scaling, wrap) (SDFChainRadix22.scala:487-493); ifft = swap re/im in and out
    (:205-211,362-383)"""
    rng = np.random.default_rng(6)
    n, log_n = 64, 6
    re, im = rng.integers(-100, 100, n), rng.integers(-100, 100, n)
    cfg = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, 0, True)       # never scale: plain FFT while no overflow
    ore, oim = oracle.fft_fixed(cfg, re, im)
    want = np.fft.fft(re + 1j * im)
    assert np.abs(ore - want.real).max() <= 6 and np.abs(oim - want.imag).max() <= 6   # 2 twiddle roundings, unscaled
    inv = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, (1 << log_n) - 1, False)
    br, bi = oracle.fft_fixed(inv, ore, oim)                                          # scaled by 1/N -> recovers the input
    assert np.abs(br - re).max() <= 3 and np.abs(bi - im).max() <= 3


# ------------------------------------------------------------------------------------------------ radix-2 chain (SDFChainRadix2.scala)
def real_tone(n, f1r):   # Radix2Spec.getRealTone (Radix2Spec.scala:18-21), f2r = 0
    i = np.arange(n)
    return np.sin(2 * np.pi * f1r * i) + np.sin(2 * np.pi * 0.0 * i)


@pytest.mark.parametrize("dit", [False, True])
@pytest.mark.parametrize("n", [4, 8, 16, 32, 64, 128, 256, 512])
def test_fft_radix2_spec_tolerance(oracle, n, dit):
    """Radix2Spec.scala:27-52: radix 2, DIF and DIT, Convergent, scale every stage, real tone at 1/N, testFFT(3, ...) with the
    default FFTParams.fixed widths (16-bit Q2.14, SDFFFTParams.scala:92-96): |raw error| <= 2^3 - 1 vs fft(x) / N;
    the 4-multiplier variant (:55-70) gives the same bits because the product is taken with NoTrim"""
    log_n = n.bit_length() - 1
    x = real_tone(n, 1.0 / n)
    want = np.fft.fft(x) / n * (1 << 14)
    outs = []
    for use4 in (False, True):
        cfg = oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", use4, True, (1 << log_n) - 1, True, "2", dit)
        ore, oim = oracle.fft_fixed(cfg, quantize(x, 14), np.zeros(n, np.int64))
        err = max(np.abs(ore - want.real).max(), np.abs(oim - want.imag).max())
        assert err <= 7, f"N={n} dit={dit}: max raw error {err:.2f} LSB"
        outs.append((ore, oim))
    assert (outs[0][0] == outs[1][0]).all() and (outs[0][1] == outs[1][1]).all()


@pytest.mark.parametrize("n", [2, 4, 8, 16, 32, 64, 128, 256, 512, 1024])
def test_fft_radix2_random_spec_tolerance(oracle, n):
    """Radix2RandomSpec.scala:14-40: genRandSignal, 16-bit Q2.14, Convergent, radix 2, simpleTestFFT(3, ...)"""
    log_n = n.bit_length() - 1
    for dit in (False, True):
        cfg = oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", False, True, (1 << log_n) - 1, True, "2", dit)
        for rep in range(2):
            re, im = java_rand_signal(oracle, n, 11110 + rep)
            ore, oim = oracle.fft_fixed(cfg, quantize(re, 14), quantize(im, 14))
            want = np.fft.fft(re + 1j * im) / n * (1 << 14)
            err = max(np.abs(ore - want.real).max(), np.abs(oim - want.imag).max())
            assert err <= 7, f"N={n} dit={dit}: max raw error {err:.2f} LSB"


def test_fft_radix2_orders_runtime_size_and_ifft(oracle):
    """without useBitReverse DIF leaves its bins bit-reversed and DIT expects bit-reversed input (SDFFFT.scala:217-270,
    SDFFFTTester.scala:127-129); a run-time size of k stages is a k-stage transform; ifft(fft(x)) recovers x"""
    rng = np.random.default_rng(8)
    n, log_n = 128, 7
    re, im = rng.integers(-9000, 9000, n), rng.integers(-9000, 9000, n)
    rev = np.array([int(format(i, f"0{log_n}b")[::-1], 2) for i in range(n)])
    mask = (1 << log_n) - 1
    for dit in (False, True):
        nat = oracle.fft_fixed(oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, mask, True, "2", dit), re, im)
        raw_cfg = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, False, mask, True, "2", dit)
        if dit:
            raw = oracle.fft_fixed(raw_cfg, re[rev], im[rev])          # bit-reversed in, natural out
            assert (raw[0] == nat[0]).all() and (raw[1] == nat[1]).all()
        else:
            raw = oracle.fft_fixed(raw_cfg, re, im)                    # natural in, bit-reversed out
            assert (raw[0] == nat[0][rev]).all() and (raw[1] == nat[1][rev]).all()
        want = np.fft.fft(re + 1j * im) / n
        assert max(np.abs(nat[0] - want.real).max(), np.abs(nat[1] - want.imag).max()) <= 4
        sre, sim = re // 100, im // 100                                # small enough for a transform that never scales
        fwd = oracle.fft_fixed(oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, 0, True, "2", dit), sre, sim)
        inv = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, mask, False, "2", dit)   # scaled inverse: 1/N
        br, bi = oracle.fft_fixed(inv, fwd[0], fwd[1])
        assert max(np.abs(br - sre).max(), np.abs(bi - sim).max()) <= 3


@pytest.mark.parametrize("n", [4, 8, 16, 32, 64, 128, 256, 512, 1024])
def test_fft_radix22_dit_tolerance(oracle, n):
    """radix 2^2 with DITDecimType (Radix22Spec.scala:31-85 runs both decimation types with testFFT(3, ...) on a real tone;
    Radix22RandomSpec's recipe added): within t = 3 of fft(x) / N, for an even and an odd elaborated stage count (a run-time size of the
    first k stages of a chain one stage longer), and equal to DIF within the rounding noise"""
    log_n = n.bit_length() - 1
    mask = (1 << log_n) - 1
    x = real_tone(n, 1.0 / n)
    re, im = java_rand_signal(oracle, n, 11110)
    for smax in (log_n, log_n + 1):
        cfg = oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", False, True, mask, True, "2^2", True, smax)
        for sr, si in ((x, np.zeros(n)), (re, im)):
            ore, oim = oracle.fft_fixed(cfg, quantize(sr, 14), quantize(si, 14))
            want = np.fft.fft(sr + 1j * si) / n * (1 << 14)
            err = max(np.abs(ore - want.real).max(), np.abs(oim - want.imag).max())
            assert err <= 7, f"N={n} numStages={smax}: max raw error {err:.2f} LSB"
    dif = oracle.fft_fixed(oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", False, True, mask, True), quantize(re, 14), quantize(im, 14))
    dit = oracle.fft_fixed(oracle.fft_cfg(log_n, 16, 14, 16, "Convergent", False, True, mask, True, "2^2", True), quantize(re, 14), quantize(im, 14))
    assert max(np.abs(dif[0] - dit[0]).max(), np.abs(dif[1] - dit[1]).max()) <= 6


def test_radix2_rom_is_run_0_of_the_radix22_table(oracle):
    """W_(2 delay)^k (SDFChainRadix2.scala:300-306) equals W_(4 delay)^(2k) (SDFChainRadix22.scala:262-267) bit for bit: the CUDA
    library reads the radix-2 stage ROMs out of its radix-2^2 tables (rsp_host.hpp build_tables)"""
    for tw in (16, 12, 9):
        for d in range(1, 15):
            delay = 1 << d
            r2 = oracle.twiddle_table_radix2(delay, tw)
            r4 = oracle.twiddle_table(4 * delay, tw)
            assert (r2[0] == r4[0][0:2 * delay:2]).all() and (r2[1] == r4[1][0:2 * delay:2]).all()


def test_trim_modes(oracle):
    """dsptools TrimType semantics on x / 2 (names of Radix22SQNRSpec.scala:184)"""
    table = {  # value: (-3, -1, 1, 3) / 2 = (-1.5, -0.5, 0.5, 1.5)
        "RoundHalfUp": (-1, 0, 1, 2), "RoundDown": (-2, -1, 0, 1), "Convergent": (-2, 0, 0, 2), "RoundUp": (-1, 0, 1, 2),
        "RoundTowardsZero": (-1, 0, 0, 1), "RoundTowardsInfinity": (-2, -1, 1, 2), "RoundHalfDown": (-2, -1, 0, 1),
        "RoundHalfTowardsZero": (-1, 0, 0, 1), "RoundHalfTowardsInfinity": (-2, -1, 1, 2), "RoundHalfToOdd": (-1, -1, 1, 1),
    }
    for mode, want in table.items():
        assert tuple(oracle.trim(v, 1, mode) for v in (-3, -1, 1, 3)) == want, mode
    assert oracle.trim(5, 2, "RoundHalfUp") == 1 and oracle.trim(6, 2, "RoundHalfUp") == 2 and oracle.trim(-6, 2, "RoundHalfUp") == -1
    assert oracle.trim(7, 1, "NoTrim") == 7


def test_twiddle_table(oracle):
    """(cos, -sin)(2 pi t / stageN) in Q2.14 (SDFChainRadix22.scala:262-267)"""
    re, im = oracle.twiddle_table(64, 16)
    t = np.arange(48)
    assert (re == np.floor(np.cos(2 * np.pi * t / 64) * 16384 + 0.5)).all()
    assert (im == np.floor(-np.sin(2 * np.pi * t / 64) * 16384 + 0.5)).all()
    assert re[0] == 16384 and im[16] == -16384


# ------------------------------------------------------------------------------------------------- LogMagMux

def test_mag_closed_forms_integer_inputs(oracle):
    """LogMagMuxSpec.scala:47-66 ("binaryPoint set to zero"): 8-bit-range integers, JPL and |z|^2 exact for integers"""
    r = oracle.JavaRandom(11110)
    rng = np.random.default_rng(11110)
    re = rng.integers(-128, 128, 200)
    im = rng.integers(-128, 128, 200)
    words = oracle.pack_iq(re, im)
    for sel, tol in ((2, 3), (1, 7)):
        cfg = oracle.mag_cfg(16, 0, 16, 11, 11, sel)
        out = oracle.mag_words(cfg, words)
        hi, lo = oracle.unpack_iq(out)
        assert (hi == 0).all()                                # .real := 0 (LogMagMux.scala:147-163)
        u, v = np.maximum(abs(re), abs(im)), np.minimum(abs(re), abs(im))
        want = np.maximum(u + v / 8, 7 * u / 8 + v / 2) if sel == 2 else re * re + im * im
        assert np.abs(lo - want).max() <= (1 << tol) - 1
        if sel == 1:
            assert (lo == want).all()                         # integer products are exact at binPoint 0
    out = oracle.mag_words(oracle.mag_cfg(16, 0, 16, 11, 11, 0), words)     # bypass
    assert (out == words).all()


def test_mag_fractional_inputs_tolerance(oracle):
    """LogMagMuxSpec.scala:21-31: inputs (U*2-1)*50 in Q(16,8)-like precision, tol t = 3 (JPL) / 7 (|z|^2)"""
    rng = np.random.default_rng(1)
    bp = 4                                                    # 16-bit, 4 fractional bits: +-50 and 2*50^2 = 5000 fit
    re, im = (rng.uniform(-1, 1, 300) * 50), (rng.uniform(-1, 1, 300) * 50)
    qr, qi = np.floor(re * 16 + 0.5).astype(np.int64), np.floor(im * 16 + 0.5).astype(np.int64)
    rq, iq = qr / 16.0, qi / 16.0
    for sel, tol in ((2, 3), (1, 7)):
        out = oracle.mag_words(oracle.mag_cfg(16, bp, 16, 9, 9, sel), oracle.pack_iq(qr, qi))
        _, lo = oracle.unpack_iq(out)
        want = np.array([oracle.jpl_f64(a, b) for a, b in zip(rq, iq)]) if sel == 2 else (rq * rq + iq * iq) / 16.0 * 16.0
        if sel == 1:
            lo = lo.astype(np.int64) & 0xFFFF                 # 5000 * 16 = 80000 wraps the 16-bit output (the reference notes
            want_raw = np.floor(want * 16).astype(np.int64) & 0xFFFF   # "scale to prevent overflow"); compare modulo 2^16
            d = (lo - want_raw + 32768) % 65536 - 32768
            assert np.abs(d).max() <= (1 << tol) - 1
        else:
            assert np.abs(lo - want * 16).max() <= (1 << tol) - 1


def test_log2_lut_and_intended_math(oracle):
    """sel 3: out = (floor(log2 jpl) - bp) << bpLog + LUT[fraction bits]  (LogMagMux.scala:115-143); vs log2(jpl) within
    2^-L + LUT rounding.  The reference never checks log2 (LogMagMuxTester.scala:64-66): unpinned."""
    tab = oracle.log2_table(9, 9)
    n = np.arange(512)
    assert (tab == np.floor(np.log2(1 + n / 512.0) * 512 + 0.5)).all()
    rng = np.random.default_rng(2)
    re, im = rng.integers(1, 20000, 500), rng.integers(-20000, 20000, 500)
    out = oracle.mag_words(oracle.mag_cfg(16, 0, 16, 9, 9, 3), oracle.pack_iq(re, im))
    _, lo = oracle.unpack_iq(out)
    jpl = np.array([np.floor(oracle.jpl_f64(a, b) + 1e-9) for a, b in zip(re, im)])
    u, v = np.maximum(abs(re), abs(im)), np.minimum(abs(re), abs(im))
    jpl = np.maximum(u + (v >> 3), (u - (u >> 3)) + (v >> 1))
    assert np.abs(lo / 512.0 - np.log2(jpl)).max() < 2.0 ** -8


def test_window_rom_and_multiply():
    """window ROM (WindowFunc.scala:41-65 periodic forms, protoWin = FixedPoint(w, w - 2), SDFFFTParams.scala:118) and the
    truncating multiply of SDFFFT.scala:233-234"""
    from oracle import oracle as O
    n = 256
    han = O.window_coeffs(O.WINDOW_HANNING, n)
    ham = O.window_coeffs(O.WINDOW_HAMMING, n, 0.54, 0.46)
    bla = O.window_coeffs(O.WINDOW_BLACKMAN, n, 0.42, 0.5, 0.08)
    tri = O.window_coeffs(O.WINDOW_TRIANGULAR, n)
    for w in (han, ham, bla, tri):
        assert w.shape == (n,) and np.allclose(w[1:], w[1:][::-1]) and abs(w[n // 2] - 1.0) < 1e-12     # periodic: w[k] = w[n - k]
    assert han[0] == 0.0 and abs(ham[0] - 0.08) < 1e-12 and abs(bla[0]) < 1e-12 and tri[0] == 0.0 and tri[n // 4] == 0.5
    assert O.window_coeffs(O.WINDOW_HANNING, 1)[0] == 1.0
    raw = O.window_raw(han, 16)
    assert raw[n // 2] == 1 << 14 and raw[0] == 0 and raw.max() == 1 << 14 and (raw == np.floor(han * 2 ** 14 + 0.5)).all()
    assert O.window_raw([1.0, -1.0, 1.99], 12).tolist() == [1 << 10, -(1 << 10), 2038]
    # all ones (WindowFunctionTypes.None) is the identity, also on the -32768 corner
    rng = np.random.default_rng(0)
    beats = rng.integers(0, 1 << 32, (3, n), dtype=np.uint64).astype(np.uint32)
    beats[0, 0] = 0x80008000
    assert (O.apply_window_words(beats, O.window_raw(np.ones(n), 16)) == beats).all()
    # floor, not round: -1 * 0.5 -> -1 (arithmetic shift), 1 * 0.5 -> 0
    b = np.array([[(0xFFFF << 16) | 0x0001]], dtype=np.uint32)
    assert O.apply_window_words(b, O.window_raw([0.5], 16))[0, 0] == 0xFFFF0000
    # run-time size below numPoints: the first n ROM entries
    assert (O.apply_window_words(beats[:, :64], raw) == O.apply_window_words(np.pad(beats[:, :64], ((0, 0), (0, n - 64))), raw)[:, :64]).all()


# ---------------------------------------------------------------------------------------------------- wide data, grow stages, overflow flags
def pack_beats64(re, im, w):
    m = (1 << w) - 1
    return ((np.asarray(re, np.int64).astype(np.uint64) & np.uint64(m)) << np.uint64(w)) | (np.asarray(im, np.int64).astype(np.uint64) & np.uint64(m))


def unpack_beats64(words, w):
    """Cat(re, im) with w-bit halves out of 64-bit beats [..., n] or 128-bit beats [..., n, 2] (low word first)"""
    words = np.asarray(words, np.uint64)
    if words.ndim >= 2 and words.shape[-1] == 2 and 2 * w > 64:
        lo, hi = words[..., 0].astype(object), words[..., 1].astype(object)
        full = lo + (hi << 64)
    else:
        full = words.astype(object)
    m = (1 << w) - 1
    def sx(v):
        v = v & m
        return v - (1 << w) if v >> (w - 1) else v
    f = np.vectorize(lambda v: sx(int(v) >> w), otypes=[object])
    g = np.vectorize(lambda v: sx(int(v)), otypes=[object])
    return f(full).astype(np.float64), g(full).astype(np.float64)


@pytest.mark.parametrize("name,dw,bp,grow,want_db", [("24-bit scaled", 24, 22, False, 98), ("16-bit grow", 16, 14, True, 82), ("24-bit grow", 24, 22, True, 130)])
def test_fft_reproduces_published_sqnr_of_wide_and_grow_configs(oracle, name, dw, bp, grow, want_db):
    """Radix22SQNRSpec.scala:60-150 at N = 1024, Convergent, twiddleWidth = dataWidth: the published chart (SQNRdataWidth.png) reads
    ~98 dB (24-bit, scale every stage), ~82 dB (16-bit, expandLogic all ones), ~130 dB (24-bit, expandLogic all ones); +-3 dB.
    Grow stages keep the butterfly outputs in a type one bit wider (SDFFFTParams.scala:119-123): the result is FFT(x), not FFT(x)/N,
    and a 24-bit input comes out 34 bits wide (128-bit beats)."""
    log_n, n = 10, 1024
    cfg = oracle.fft_cfg(log_n, dw, bp, dw, "Convergent", False, True, (1 << log_n) - 1, True, expand_mask=(1 << log_n) - 1 if grow else 0)
    wo = oracle.fft_stage_width(cfg, log_n - 1)
    assert wo == dw + (log_n if grow else 0)
    vals = []
    for rep in range(6):
        re, im = java_rand_signal(oracle, n, 300 + rep)
        out, ovf = oracle.fft_fixed_words64(cfg, pack_beats64(quantize(re, bp), quantize(im, bp), dw)[None, :])
        ore, oim = unpack_beats64(out[0], wo)
        ref = np.fft.fft(re + 1j * im) / (1 if grow else n)
        vals.append(sqnr_db((ore + 1j * oim) / (1 << bp), ref))
        if grow:
            assert ovf == 0          # a grow stage cannot overflow (SDFChainRadix22.scala:455-457)
    assert abs(np.mean(vals) - want_db) < 3.0, f"{name}: {np.mean(vals):.1f} dB vs published ~{want_db}"


def test_fft_overflow_flags(oracle):
    """per-stage overflow status (SDFChainRadix22.scala:473-486): `tooBig` = a grown butterfly output at or above 2^(w-1); the check
    looks at the sum BEFORE div2, so it also fires on stages that scale.  Hand-derived: a constant row x = (A, 0) doubles its DC term
    at every stage's sum before the halving: flag at stage 0 iff 2A >= 2^15; a row below half scale never flags; full-scale square
    wave flags the stages where it adds up."""
    log_n, n = 6, 64
    cfg = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, (1 << log_n) - 1, True)
    for amp, want0 in ((16383, 0), (16384, 1)):
        _, _, m = oracle.fft_fixed_ex(cfg, np.full(n, amp), np.zeros(n, np.int64))
        # the sum outputs keep the value (A + A) / 2 = A: the same test repeats at every stage, on the DC path
        assert m == (((1 << log_n) - 1) if want0 else 0)
    # difference path: +A, -A alternating at distance N/2 -> stage 0 difference 2A
    x = np.where(np.arange(n) < n // 2, 20000, -20000)
    _, _, m = oracle.fft_fixed_ex(cfg, x, np.zeros(n, np.int64))
    assert m & 1
    # keep-LSB stages (no scaling) report the same flag, and there the value really wraps
    cfg2 = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, 0, True)
    ore, _, m2 = oracle.fft_fixed_ex(cfg2, np.full(n, 600), np.zeros(n, np.int64))
    assert ore[0] == ((600 * 64 + 32768) % 65536) - 32768 and m2 == 0b100000     # 600 * 2^6 = 38400 wraps at the last stage only
    # radix-2 chain and DIT carry the same check (SDFChainRadix2.scala:409-423)
    for radix, dit in (("2", False), ("2", True), ("2^2", True)):
        c = oracle.fft_cfg(log_n, 16, 0, 16, "RoundHalfUp", False, True, (1 << log_n) - 1, True, sdf_radix=radix, decim_dit=dit, num_stages=log_n)
        _, _, m = oracle.fft_fixed_ex(c, np.full(n, 16384), np.zeros(n, np.int64))
        assert m != 0
        _, _, m = oracle.fft_fixed_ex(c, np.full(n, 100), np.zeros(n, np.int64))
        assert m == 0
