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
    N, TPC = 1 << S, (1 << S) // 16
    assert sum(pass_size(S, k) for k in range(n_passes(S))) == S
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        assert i0 % 2 == 0                       # a radix-2^2 pair (butterfly, -j, butterfly, twiddle) never straddles two passes
        LDl = S - i0 - m; Dl = 1 << LDl; G = 16 >> m; R = 1 << m
        for tl in range(TPC):
            base0 = ((tl >> LDl) << (m + LDl)) | (tl & (Dl - 1))
            for g in range(G):
                vt = g * TPC + tl
                base = ((vt >> LDl) << (m + LDl)) | (vt & (Dl - 1))
                assert base == base0 + g * TPC * R
                for rr in range(R):
                    assert xidx(base + rr * Dl, S) == xidx(base0, S) + xidx(g * TPC * R + rr * Dl, S)
            assert midx(16 * tl + 15) == midx(16 * tl) + 15
        assert xidx(N - 1, S) < N + (N >> 3) + (N >> 4) + 8


@pytest.mark.parametrize("S,worst", [(10, 1.0), (11, 1.0), (12, 1.0), (13, 1.25), (14, 1.0)])
def test_exchange_buffer_padding_is_conflict_free(S, worst):
    N, TPC = 1 << S, (1 << S) // 16
    tot = ideal = 0
    for K in range(n_passes(S)):
        m, i0 = pass_size(S, K), pass_start(S, K)
        LDl = S - i0 - m; Dl = 1 << LDl; G = 16 >> m; R = 1 << m
        last = K == n_passes(S) - 1
        for warp0 in range(0, min(TPC, 64), 32):
            for g in range(G):
                for rr in range(R):
                    addrs = []
                    for lane in range(32):
                        vt = g * TPC + (warp0 + lane) % TPC
                        base = ((vt >> LDl) << (m + LDl)) | (vt & (Dl - 1))
                        addrs.append(xidx(base + rr * Dl, S) * 8)
                    if last and rr % 2:
                        continue                     # the last pass reads two neighbouring points per 128-bit load
                    w, i = wavefronts(addrs, 16 if last else 8)
                    tot += w; ideal += i
    assert tot / ideal <= worst + 1e-9
