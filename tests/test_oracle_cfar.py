"""Pins the oracle's CFAR restatement:
 - the Scala golden models CFARUtils.cfarCA / cfarOS / cfarGOS / cfarCASH (generators/cfar/src/main/scala/CFARUtils.scala:29-272)
   restated in C (oracle/rsp_oracle.c) are checked against an independent numpy restatement written here from the same source
 - the integer hardware cores are held to the reference's own acceptance (CFARCATester.scala:120-131): thresholds within
   fixTolLSBs t = 3 (|raw err| <= 7) of the golden in Q10.6 with scaler 3.5, on the reference's seeded test signal
   (|FFT| of 3 tones + noise, Random.setSeed(11110L), CFARCATester.scala:231-244), and the golden's peaks all flagged."""
import numpy as np
import pytest

MODES = ["Cell Averaging", "Greatest Of", "Smallest Of"]


def reference_cfar_signal(oracle, n):
    """CFARCATester.scala:231-244"""
    r = oracle.JavaRandom(11110)
    noise = np.array([np.sqrt(r.next_double() + r.next_double()) for _ in range(n)])
    i = np.arange(n)
    s = noise + 0.4 * np.exp(2j * np.pi * i / 8) + 0.2 * np.exp(2j * np.pi * i / 4) + 0.1 * np.exp(2j * np.pi * i / 2)
    return np.abs(np.fft.fft(s))


def np_cfar_ca(sig, mode, R, G, edges, scale):
    """independent restatement of cfarCA (CFARUtils.scala:84-131)"""
    n, wc = len(sig), R + G
    comb = {"Cell Averaging": lambda a, b: (a + b) / 2, "Greatest Of": max, "Smallest Of": min}[mode]
    thr, peaks = [], []
    if edges:
        for ix in range(n):
            lead = 0.0 if ix < wc else sig[ix - wc:ix - G].sum() / R
            lagg = 0.0 if ix >= n - wc else sig[ix + G + 1:ix + wc + 1].sum() / R
            t = lagg if ix < wc else (lead if ix >= n - wc else comb(lead, lagg))
            thr.append(t * scale)
        return np.array(thr), peaks
    for ix in range(wc, n - wc):
        t = comb(sig[ix - wc:ix - G].sum() / R, sig[ix + G + 1:ix + wc + 1].sum() / R) * scale
        if t < sig[ix]:
            peaks.append(ix)
        thr.append(t)
    return np.array([0.0] * wc + thr + [0.0] * wc), peaks


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("R,G", [(16, 2), (16, 4), (32, 4), (64, 8)])
def test_golden_ca_matches_independent_restatement(oracle, mode, R, G):
    sig = reference_cfar_signal(oracle, 512)
    for edges in (False, True):
        thr, peaks = oracle.golden_cfar_ca(sig, mode, R, G, edges, 3.5)
        t2, p2 = np_cfar_ca(sig, mode, R, G, edges, 3.5)
        assert thr.shape == t2.shape and np.allclose(thr, t2, rtol=1e-12, atol=1e-9)
        assert peaks.tolist() == p2


def test_golden_os_gos_cash_against_numpy(oracle):
    sig = reference_cfar_signal(oracle, 256)
    n, R, G = 256, 16, 3
    wc = R + G
    # cfarOS: ascending sort of lead ++ lagg, element (1.5 * R).toInt; peaks appended twice (CFARUtils.scala:165-174)
    thr, peaks = oracle.golden_cfar_os(sig, R, G, 2.0)
    want = [0.0] * wc + [np.sort(np.concatenate([sig[ix - wc:ix - G], sig[ix + G + 1:ix + wc + 1]]))[int(1.5 * R)] * 2.0
                         for ix in range(wc, n - wc)] + [0.0] * wc
    assert np.allclose(thr, want)
    pk = [ix for ix in range(wc, n - wc) if want[ix] < sig[ix]]
    assert peaks.tolist() == [p for p in pk for _ in (0, 1)]
    # cfarGOS, considerEdges = false (CFARUtils.scala:247-262)
    thr, peaks = oracle.golden_cfar_gos(sig, R, G, 12, 5, "Greatest Of", False, 2.0)
    want = [0.0] * wc + [max(np.sort(sig[ix - wc:ix - G])[::-1][5 - 1], np.sort(sig[ix + G + 1:ix + wc + 1])[::-1][12 - 1]) * 2.0
                         for ix in range(wc, n - wc)] + [0.0] * wc
    assert np.allclose(thr, want)
    # cfarCASH (CFARUtils.scala:29-81), guardCells = 0 as the reference tester calls it (CFARCATester.scala:93)
    sub = 4
    thr, peaks = oracle.golden_cfar_cash(sig, R, sub, 0, 3.5)
    ns = R // sub
    want = []
    for ix in range(R, n - R):
        mx = []
        for s in range(ns, 0, -1):
            lead = sig[ix - s * sub:ix - (s - 1) * sub].sum() / sub
            lagg = sig[ix + (ns - s) * sub + 1:ix + (ns - s + 1) * sub + 1].sum() / sub
            mx.append(max(lead, lagg))
        want.append(min(mx) * 3.5)
    assert np.allclose(thr, [0.0] * R + want + [0.0] * R)
    assert peaks.tolist() == [ix for ix, t in zip(range(R, n - R), want) if t < sig[ix]]


@pytest.mark.parametrize("mode_num,mode", list(enumerate(MODES)))
@pytest.mark.parametrize("W", [16, 32, 64])
@pytest.mark.parametrize("G", [2, 4, 8])
def test_mem_core_meets_reference_acceptance(oracle, mode_num, mode, W, G):
    """CFARCoreWithMem in Q10.6 (CFARSpecFinal.scala: protoIn FixedPoint(16, 6)), scaler 3.5, divSum = log2 W
    (CFARCATester.scala:106): thresholds within t = 3 of cfarCA(considerEdges = true); identical detections inside"""
    n, bp = 512, 6
    sig = reference_cfar_signal(oracle, n)
    x = np.floor(sig * (1 << bp) + 0.5).astype(np.int64)
    assert x.max() < (1 << 15)
    cfg = oracle.cfar_cfg(n, W, G, int(3.5 * (1 << bp)), core=oracle.CORE_MEM, in_bp=bp, thr_width=24, thr_bp=bp + 1,
                          scaler_bp=bp, log_or_linear=1, cfar_mode=mode_num, lead_lagg_window_size=64, guard_window_size=8)
    thr_full, thr, peak = oracle.cfar_fixed(cfg, x)
    g_thr, _ = oracle.golden_cfar_ca(sig, mode, W, G, True, 3.5)
    err = thr / float(1 << (bp + 1)) - g_thr
    assert np.abs(err * (1 << bp)).max() <= 7.0, f"max threshold error {np.abs(err * (1 << bp)).max():.2f} LSB"
    _, g_peaks = oracle.golden_cfar_ca(sig, mode, W, G, False, 3.5)
    wc = W + G
    inner = np.nonzero(peak[wc:n - wc])[0] + wc
    assert set(g_peaks.tolist()) <= set(inner.tolist())            # the one-way check of CFARCATester.scala:129-131
    assert set(inner.tolist()) == set(g_peaks.tolist())            # and (stronger) identical detection sets


def test_lis_and_asr_cores_against_goldens(oracle):
    """GOS (CFARCoreWithLis) vs cfarGOS and CASH (CFARCoreWithASR) vs cfarCASH: exact on integer-valued signals at
    binPoint 0 with scaler 1 where no rounding occurs in the order statistics; edges are 0 (CFARCoreWithLis.scala:172-173)"""
    rng = np.random.default_rng(7)
    n, W, G = 256, 16, 3
    x = rng.integers(0, 4000, n)
    for mode_num, mode in enumerate(MODES):
        cfg = oracle.cfar_cfg(n, W, G, 2, core=oracle.CORE_LIS, cfar_algorithm=1, cfar_mode=mode_num, index_lagg=12, index_lead=5)
        _, thr, peak = oracle.cfar_fixed(cfg, x)
        g_thr, g_peaks = oracle.golden_cfar_gos(x.astype(float), W, G, 12, 5, mode, False, 2.0)
        want = np.floor(g_thr) if mode_num == 0 else g_thr       # CA halves with >> 1 (floor) before scaling
        if mode_num == 0:
            want = np.array([0.0] * (W + G) + [2.0 * ((np.sort(x[ix - W - G:ix - G])[::-1][4] + np.sort(x[ix + G + 1:ix + G + W + 1])[::-1][11]) >> 1)
                                               for ix in range(W + G, n - W - G)] + [0.0] * (W + G))
        assert (thr == want).all()
        if mode_num != 0:
            inner = np.nonzero(peak[W + G:n - W - G])[0] + W + G
            assert inner.tolist() == sorted(set(g_peaks.tolist()))
    # CASH: sub-window sums, min of max, >> log2(subCells)
    sub = 4
    cfg = oracle.cfar_cfg(n, W, G, 3, core=oracle.CORE_ASR, cfar_mode=3, sub_cells=sub, div_sum=2)
    _, thr, _ = oracle.cfar_fixed(cfg, x)
    for ix in (W + G, 100, n - W - G - 1):
        mx = [max(x[ix + G + W - i * sub - sub + 1:ix + G + W - i * sub + 1].sum(), x[ix - G - i * sub - sub:ix - G - i * sub].sum())
              for i in range(W // sub)]
        assert thr[ix] == 3 * (min(mx) >> 2)
    assert thr[0] == 0 and thr[n - 1] == 0


def test_cfar_word_packing_and_edges(oracle):
    """Cat(threshold, fftBin, peak) (CFARDspBlock.scala:90); one-sided windows at the edges (CFARCoreWithMem.scala:154-171)"""
    n, W, G = 64, 8, 2
    x = np.arange(n) * 10
    cfg = oracle.cfar_cfg(n, W, G, 1, core=oracle.CORE_MEM, cfar_mode=0, div_sum=3)
    _, thr, peak = oracle.cfar_fixed(cfg, x)
    assert thr[0] == x[G + 1:G + W + 1].sum() >> 3                # lagging window only
    assert thr[n - 1] == x[n - 1 - G - W:n - 1 - G].sum() >> 3    # leading window only
    mid = 30
    assert thr[mid] == ((x[mid + G + 1:mid + G + W + 1].sum() >> 3) + (x[mid - G - W:mid - G].sum() >> 3)) >> 1
    words = oracle.cfar_words(cfg, x.astype(np.uint32))
    assert ((words >> 1) & 63 == np.arange(n)).all()
    assert ((words & 1) == peak).all()
    assert ((words >> 7) == (thr & 0xFFFF)).all()
    with pytest.raises(ValueError):                               # assert(fftWin > 2*windowCells + 2*guardCells + 1)
        oracle.cfar_fixed(oracle.cfar_cfg(16, 8, 2, 1), np.zeros(16))


@pytest.mark.parametrize("R,G,scaler,log_mode", [(32, 4, 2, False), (16, 1, 3, False), (64, 8, 1, False), (32, 4, 700, True), (8, 0, 5, False)])
def test_integer_os_core_equals_the_golden_cfaros(oracle, R, G, scaler, log_mode):
    """BASELINE configs[3] reading (B): the integer combined-window core (core = CORE_OS) against the Scala golden cfarOS
    (CFARUtils.scala:159-196) on integer-valued profiles (exponential clutter + targets, the C4 recipe of SURVEY 8d): identical
    thresholds everywhere (zero edges included) and the golden's peak set (each peak listed twice by the golden, :172,174)"""
    rng = np.random.default_rng(R + G)
    n = 4096
    x = np.floor(rng.exponential(256.0, n)).astype(np.int32)
    x[[300, 1200, 2500, 3900]] += 4096
    x[1000:1040] = 77          # a plateau of ties
    cfg = oracle.cfar_cfg(n, R, G, scaler, core=oracle.CORE_OS, log_or_linear=0 if log_mode else 1, thr_width=32, in_width=16,
                          scaler_width=16, guard_window_size=max(G, 1))
    thr_full, thr, peak = oracle.cfar_fixed(cfg, x)
    g_thr, g_peaks = oracle.golden_cfar_os(x.astype(np.float64), R, G, float(scaler), log_mode)
    assert g_thr.shape == (n,)
    assert (thr_full == g_thr.astype(np.int64)).all()
    assert sorted(set(g_peaks.tolist())) == np.flatnonzero(peak).tolist()
    assert len(g_peaks) == 2 * int(peak.sum()) and peak.sum() >= 4
    # the hardware reading (A) of the same words "ref 64, k = 48" is a different statistic: both labelled in bench.py
    if R == 32 and not log_mode:
        cfg_a = oracle.cfar_cfg(n, 64, G, scaler, core=oracle.CORE_LIS, cfar_algorithm=1, index_lagg=48, index_lead=48, thr_width=32)
        thr_a, _, _ = oracle.cfar_fixed(cfg_a, x)
        assert (thr_a != thr_full).any()
