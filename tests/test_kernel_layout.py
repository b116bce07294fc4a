"""CPU model of the shared-memory layout of the fp32 fast kernel (rsp_chains_b200/csrc/rsp_fast_f32.cu): the kernel addresses the
exchange buffer as ONE run-time base plus compile-time offsets, which is only valid if xidx is additive for the shapes used, and it
relies on the padding p + 2 (p >> 4) + 4 (p >> 6) (no second term at N = 4096) being bank-conflict free for every pass.  Both are checked exhaustively here, so a
change of the pass decomposition (rsp_device.cuh, fx_pass_*) or of the padding that breaks either shows up without a GPU."""
import pytest


def n_passes(S): return 1 if S <= 4 else (S // 4 if S % 4 == 0 else S // 4 + 1)


def pass_size(S, k):
    if S <= 4: return S
    if S % 4 == 0: return 4
    if S % 4 == 2: return 4 if k < S // 4 else 2
    if S % 4 == 3: return 4 if k < S // 4 else 3
    return 4 if k < S // 4 - 1 else (2 if k == S // 4 - 1 else 3)


def pass_start(S, k): return 0 if k == 0 else pass_start(S, k - 1) + pass_size(S, k - 1)


def xidx(p, S=10): return p + ((p >> 4) << 1) + (0 if S == 12 else ((p >> 6) << 2))


def midx(n): return n + ((n >> 5) << 2)


def _rev(x, bits):
    r = 0
    for i in range(bits): r |= ((x >> i) & 1) << (bits - 1 - i)
    return r


def f32_pts(S): return 32 if S >= 14 else 16     # FGeoF<S>::PTS: two sets of 16 points per thread from RSP_F32_NQ2_SMIN on


def gbase(v, m, LDl): return ((v >> LDl) << (m + LDl)) | (v & ((1 << LDl) - 1))


def wavefronts(addrs, width):
    """shared-memory wavefronts of one warp access: `width` bytes per lane, 128 bytes per wavefront group"""
    group, tot = 128 // width, 0
    for g0 in range(0, 32, group):
        banks = {}
        for a in addrs[g0:g0 + group]:
            for w in range(width // 4):
                banks.setdefault((a // 4 + w) % 32, set()).add(a // 4 + w)
        tot += max(len(v) for v in banks.values())
    return tot, 32 // group


@pytest.mark.parametrize("S", [10, 11, 12, 13, 14])
def test_exchange_buffer_offsets_are_additive_and_passes_start_even(S):
    PTS = f32_pts(S)
    N, TPC = 1 << S, (1 << S) // PTS
    assert sum(pass_size(S, k) for k in range(n_passes(S))) == S
    seen = set()
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        assert i0 % 2 == 0                       # a radix-2^2 pair (butterfly, -j, butterfly, twiddle) never straddles two passes
        LDl = S - i0 - m; Dl = 1 << LDl; G = PTS >> m; R = 1 << m
        for tl in range(TPC):
            base0 = gbase(tl, m, LDl)
            for g in range(G):
                vt = g * TPC + tl
                base = gbase(vt, m, LDl)
                assert base == base0 + gbase(g * TPC, m, LDl)
                assert (vt & (Dl - 1)) == (tl & (Dl - 1)) + ((g * TPC) & (Dl - 1))      # the twiddle index of the group
                for rr in range(R):
                    assert xidx(base + rr * Dl, S) == xidx(base0, S) + xidx(gbase(g * TPC, m, LDl) + rr * Dl, S)
                    if K == 0: seen.add(base + rr * Dl)
            assert midx(16 * tl + 15) == midx(16 * tl) + 15
            for h in range(PTS // 16):           # the thread's CFAR blocks of 16 bins
                assert midx(16 * (tl + h * TPC)) == midx(16 * tl) + h * (16 * TPC + ((16 * TPC) >> 3))
        assert xidx(N - 1, S) < N + (N >> 3) + (N >> 4) + 8
    assert seen == set(range(N))
    # first pass: group g, leg rr is the prefetched point pre[g + G rr] = in[tl + (g + G rr) TPC]
    G0 = PTS >> 4
    for tl in (0, 1, TPC - 1):
        for g in range(G0):
            for rr in range(16):
                assert gbase(g * TPC + tl, 4, S - 4) + rr * (N >> 4) == tl + (g + G0 * rr) * TPC
    # last pass: bin of point (vt << m) + rr = rev_m(rr) << (S - m) | rev_{S-LP}(tl) << (LP - m) | rev_{LP-m}(g)
    m = pass_size(S, n_passes(S) - 1); LP = PTS.bit_length() - 1
    for tl in (0, 1, 5, TPC - 1):
        for g in range(PTS >> m):
            for rr in range(1 << m):
                p = ((g * TPC + tl) << m) + rr
                assert _rev(p, S) == (_rev(rr, m) << (S - m)) | (_rev(tl, S - LP) << (LP - m)) | _rev(g, LP - m)


@pytest.mark.parametrize("S,worst", [(10, 1.0), (11, 1.0), (12, 1.0), (13, 1.25), (14, 1.0)])
def test_exchange_buffer_padding_is_conflict_free(S, worst):
    PTS = f32_pts(S)
    N, TPC = 1 << S, (1 << S) // PTS
    tot = ideal = 0
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        LDl = S - i0 - m; Dl = 1 << LDl; G = PTS >> m; R = 1 << m
        last = K == n_passes(S) - 1
        for warp0 in range(0, min(TPC, 64), 32):
            for g in range(G):
                for rr in range(R):
                    addrs = []
                    for lane in range(32):
                        vt = g * TPC + (warp0 + lane) % TPC
                        addrs.append(xidx(gbase(vt, m, LDl) + rr * Dl, S) * 8)
                    if last and rr % 2:
                        continue                     # the last pass reads two neighbouring points per 128-bit load
                    w, i = wavefronts(addrs, 16 if last else 8)
                    tot += w; ideal += i
    assert tot / ideal <= worst + 1e-9


# ---------------------------------------------------------------------------------------------------------------------
# fixed-point fast kernel (rsp_chains_b200/csrc/rsp_fast.cu): exchange buffer of 32-bit points, int16 magnitude array

def fidx_fixed(p, S):
    if S == 11: return p + ((p >> 5) << 2) + ((p >> 8) << 3)
    if S == 12: return p + ((p >> 5) << 2) + ((p >> 7) << 3)
    if S == 13: return p + ((p >> 5) << 3)
    return p + ((p >> 6) << 2)


def mhx(n, S): return n + ((n >> 6) << 3) + (((n >> 10) << 3) if S == 14 else 0)


def fixed_kernel_wavefronts(S):
    """(actual, ideal) shared-memory wavefronts per warp and row of the FFT exchanges, the bit-reversed magnitude scatter and the
    reading back of the magnitudes, following fast_pass / chain_fixed_fast_kernel access by access"""
    N, TPC = 1 << S, (1 << S) // 16
    got = ideal = 0
    warp = list(range(32, 64))
    def add(addrs, width):
        nonlocal got, ideal
        w, i = wavefronts(addrs, width); got += w; ideal += i
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        LDl = S - i0 - m; Dl = 1 << LDl; G = 16 >> m; R = 1 << m
        first, last = K == 0, K == n_passes(S) - 1
        for g in range(G):
            bases = [(((g * TPC + tl) >> LDl) << (m + LDl)) | ((g * TPC + tl) & (Dl - 1)) for tl in warp]
            if not first:
                if last:
                    for q in range(R // 4): add([4 * fidx_fixed(b + 4 * q, S) for b in bases], 16)
                else:
                    for rr in range(R): add([4 * fidx_fixed(b + rr * Dl, S) for b in bases], 4)
            if not last:
                for rr in range(R): add([4 * fidx_fixed(b + rr * Dl, S) for b in bases], 4)
        if last:
            for rr in range(R):
                rrr = _rev(rr, m) << (S - m)
                add([2 * mhx(rrr | (_rev(tl, S - 4) << (4 - m)), S) for tl in warp], 2 * G if G >= 2 else 4)
    for q in range(2): add([2 * mhx(16 * tl, S) + 16 * q for tl in warp], 16)
    return got, ideal


@pytest.mark.parametrize("S", [10, 11, 12, 13, 14])
def test_fixed_kernel_index_maps_are_additive(S):
    N, TPC = 1 << S, (1 << S) // 16
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        LDl = S - i0 - m; Dl = 1 << LDl; G = 16 >> m; R = 1 << m
        for tl in range(TPC):
            b0 = ((tl >> LDl) << (m + LDl)) | (tl & (Dl - 1))
            for g in range(G):
                vt = g * TPC + tl
                for rr in range(R):
                    p = (((vt >> LDl) << (m + LDl)) | (vt & (Dl - 1))) + rr * Dl
                    assert fidx_fixed(p, S) == fidx_fixed(b0, S) + fidx_fixed(g * TPC * R, S) + fidx_fixed(rr * Dl, S)
            if K == n_passes(S) - 1:      # 4 consecutive points stay contiguous and 16-byte aligned for the 128-bit loads
                for q in range(R // 4):
                    assert fidx_fixed(b0 + 4 * q, S) % 4 == 0 and fidx_fixed(b0 + 4 * q + 3, S) == fidx_fixed(b0 + 4 * q, S) + 3
    m = pass_size(S, n_passes(S) - 1)
    for tl in range(TPC):
        rtl = _rev(tl, S - 4) << (4 - m)
        for rr in range(1 << m):
            rrr = _rev(rr, m) << (S - m)
            assert mhx(rrr | rtl, S) == mhx(rrr, S) + mhx(rtl, S)
        for j in range(16): assert mhx(tl + j * TPC, S) == mhx(tl, S) + mhx(j * TPC, S)
        assert mhx(16 * tl, S) % 8 == 0 and mhx(16 * tl + 15, S) == mhx(16 * tl, S) + 15
    assert fidx_fixed(N - 1, S) + 1 >= N + 1      # the prefix sums (N + 1 words) fit the exchange buffer


@pytest.mark.parametrize("S,worst", [(10, 1.1), (11, 1.3), (12, 1.55), (13, 1.36), (14, 1.08)])
def test_fixed_kernel_bank_conflicts(S, worst):
    """the pads keep the exchanges conflict free (1.17 at N = 8192); what is left is the bit-reversed scatter of 16-bit magnitudes
    (a 4-way conflict of the 16-bit stores at N = 4096, 2-way elsewhere) -- the bound guards against a change that brings the
    16- to 32-way conflicts of the unpadded layout back"""
    got, ideal = fixed_kernel_wavefronts(S)
    assert got <= worst * ideal, (got, ideal)


# ---------------------------------------------------------------------------------------------------------------------
# FFT block alone on the fast passes (fft_fixed_fast_kernel): the bit-reversed scatter of the output beats, index wx<S>

def wx(n, S):
    return {10: n + ((n >> 5) << 2), 11: n + ((n >> 4) << 1), 12: n + (n >> 3), 13: n + ((n >> 6) << 1), 14: n + ((n >> 9) << 2)}[S]


def fixed_row_words(S):
    """FGeo<S>::ROW: exchange buffer + magnitude array, in words"""
    N = 1 << S
    xb = (fidx_fixed(N - 1, S) + 4) & ~3
    xm = (mhx(N - 1, S) // 2 + 4) & ~3
    return xb + xm


@pytest.mark.parametrize("S,LP", [(10, 4), (11, 4), (12, 4), (13, 4), (14, 4), (14, 5)])
def test_fft_only_output_scatter_is_conflict_free_and_covers_the_row(S, LP):
    m = pass_size(S, n_passes(S) - 1)
    R, G, N, TPC = 1 << m, (1 << LP) >> m, 1 << S, (1 << S) >> LP
    width = min(G, 4)                                   # words per store instruction (G = 8: two 128-bit stores)
    seen = set()
    for rr in range(R):
        rrr = _rev(rr, m) << (S - m)
        for w0 in range(0, TPC, 32):
            for part in range(max(1, G // 4)):
                addrs = []
                for lane in range(32):
                    n0 = rrr | (_rev(w0 + lane, S - LP) << (LP - m))
                    assert wx(n0, S) == wx(rrr, S) + wx(n0 - rrr, S)                 # one run-time base + a compile-time offset
                    assert wx(n0 + G - 1, S) == wx(n0, S) + G - 1 and wx(n0, S) % width == 0   # the run is contiguous and aligned
                    addrs.append(4 * (wx(n0, S) + 4 * part))
                    seen.update(range(n0, n0 + G))
                got, ideal = wavefronts(addrs, 4 * width)
                assert got == ideal, (S, LP, rr, w0)
    assert seen == set(range(N))
    assert wx(N - 1, S) < fixed_row_words(S)            # the padded output row fits the chirp's shared memory
    # reading back in natural order (bin tl + j TPC): at most two wavefronts per warp load
    for j in range(1 << LP):
        for w0 in range(0, TPC, 32):
            got, ideal = wavefronts([4 * wx(w0 + lane + j * TPC, S) for lane in range(32)], 4)
            assert got <= 2 * ideal


# ---------------------------------------------------------------------------------------------------------------------
# OS / GOS-CFAR on bit planes (chain_fixed_fast_os_kernel, chain_fixed_fast_gosb_kernel): the algorithms, step for step

def bit_transpose16(A):
    """rsp_fast.cu bit_transpose16: 8 words, word k = rows 2k | 2k+1 << 16; block swaps of size 8, 4, 2, 1"""
    A = [int(a) for a in A]
    M = 0xFFFFFFFF
    for k in range(4):
        t = ((A[k] >> 8) ^ A[k + 4]) & 0x00FF00FF; A[k + 4] ^= t; A[k] ^= (t << 8) & M
    for k in (0, 1, 4, 5):
        t = ((A[k] >> 4) ^ A[k + 2]) & 0x0F0F0F0F; A[k + 2] ^= t; A[k] ^= (t << 4) & M
    for k in (0, 2, 4, 6):
        t = ((A[k] >> 2) ^ A[k + 1]) & 0x33333333; A[k + 1] ^= t; A[k] ^= (t << 2) & M
    for k in range(8):
        t = ((A[k] >> 1) ^ (A[k] >> 16)) & 0x5555; A[k] ^= (t * 0x10002) & M
    return A


def test_bit_transpose16_turns_16_values_into_16_plane_halfwords():
    import numpy as np
    rng = np.random.default_rng(0)
    for _ in range(300):
        rows = [int(v) for v in rng.integers(0, 1 << 16, 16)]
        T = bit_transpose16([rows[2 * k] | (rows[2 * k + 1] << 16) for k in range(8)])
        for c in range(16):
            assert (T[c // 2] >> (16 * (c % 2))) & 0xFFFF == sum(((rows[r] >> c) & 1) << r for r in range(16))


def test_radix_selection_on_bit_planes_is_the_kth_largest_and_skipped_planes_change_nothing():
    """gos_select8 / os_select8 for one window: from the top plane down, t = k - popc(cand & plane); t <= 0 keeps the candidates with the bit,
    else the others and k = t.  Planes that are uniform over the row, or all zero over the window, are skipped (exact)."""
    import numpy as np
    rng = np.random.default_rng(1)
    for trial in range(400):
        W = int(rng.integers(1, 129))
        k = int(rng.integers(1, W + 1))
        hi = int(rng.choice([4, 300, 1 << 15, 1 << 16]))
        lo_bias = int(rng.choice([0, 1 << 15]))
        x = [min((int(v) + lo_bias), 0xFFFF) for v in rng.integers(0, hi, W)]      # biased (offset-binary) cells
        if trial % 7 == 0:
            x = [x[0]] * W                                                          # all ties
        row_or, row_and = 0, 0xFFFF
        for v in x: row_or |= v; row_and &= v
        act = row_or ^ row_and                                                      # planes that are not uniform
        act &= row_or                                                               # ... and not all zero over the window (os_local_planes)
        cand, kk = (1 << W) - 1, k
        for b in range(15, -1, -1):
            if not (act >> b) & 1: continue
            plane = sum(((x[i] >> b) & 1) << i for i in range(W))
            m = cand & plane
            t = kk - bin(m).count("1")
            if t <= 0: cand = m
            else: cand, kk = cand & ~plane, t
        assert cand != 0
        want = sorted(x, reverse=True)[k - 1]
        assert all(x[i] == want for i in range(W) if (cand >> i) & 1)
