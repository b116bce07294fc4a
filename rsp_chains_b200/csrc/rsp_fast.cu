// rsp_fast.cu -- the specialised fused chain kernel for the configuration the reference's chain apps run and the benchmarks
// time: 16-bit data and twiddles, binPoint 0, RoundHalfUp, scale at every stage, forward transform, natural-order output
// (src/main/scala/RspChain.scala:107-121), JPL magnitude (sel = 2, src/test/scala/FftMagCfarChainTester.scala:84), CFARCoreWithMem
// in CA / GO / SO mode with a linear or log threshold (generators/cfar/src/main/scala/CFARCoreWithMem.scala:102-189).
// Anything else runs on the generic kernels of rsp_kernels.cuh; results are identical where both apply (tests/test_gpu_chain.py).
//
// Why a second kernel: the chain is integer-issue bound, not HBM bound (DESIGN.md "kernel"), so what counts is instructions per
// sample.  Everything a register can hold at compile time is folded:
//  * butterfly (generators/sdf-fft/src/main/scala/SDFFFTUtil.scala:51-61 + div2, SDFChainRadix22.scala:459-464):
//      -s = (-a - c) >> 1 exactly, and the scaled difference is d = wrap16(s - c) because (a - c + 1) >> 1 == ((a + c + 1) >> 1) - c;
//      values are held with compile-time signs so that every add is a single IMAD on the otherwise idle FMA pipe
//  * twiddle product (SDFChainRadix22.scala:298-308): the table holds (cos, sin) << 2, so
//      (ar * cos4 - ai * sin4 + 2^15) >> 16   (32-bit wrap-around arithmetic)
//    is exactly wrap16((ar*cos - ai*sin + 2^13) >> 14): rounding, trimBinary and the 16-bit store in one shift
//  * the rows of a CTA stay in shared memory from their first global load to their last global store; the bit reversal is the
//    scatter of the magnitudes; the CFAR window sums are differences of one prefix-sum array
//  * global loads and stores are warp-contiguous 128-byte lines.
#include <algorithm>
#include <cstdint>
#include <cstdlib>

#include "rsp_fast.cuh"

namespace rsp {

// ---------------------------------------------------------------------------------------------------------------------
struct FastArgs {
  const uint32_t *in;
  uint32_t *out;
  uint32_t n_rows, row_base;
  const int32_t *win4;   // window ROM << (16 - binary point) or null (FFTParams.windowFunc, SDFFFT.scala:53-66,233-234)
  const int2 *tw;        // (c, s) with c = cos << 2, s = -sin(2 pi t / stageN) << 2; per stageN three runs [W^2k | W^k | W^3k]
  int tw_off[16];
  int W, G, div_sum, scaler;   // div_sum here = (32 - sumw) + divSum: shift applied to the pre-scaled window sums (see wmul)
  int wmul;              // 1 << (32 - sumw): the prefix sums are kept times wmul, so that differences wrap exactly like the
                         // reference's sumw-bit sum registers (CFARCoreWithMem.scala:34,108,123) in plain 32-bit arithmetic
  int one, m1;           // +1 and -1 as run-time values: multipliers of the IMADs that do the adds on the FMA pipe (see fadd)
  int rnd[4];            // 0x8000, -0x8000, 0x7FFF, -0x7FFF: rounding addends of the twiddle products (see fcmul)
  uint32_t thr_mask;     // (1 << thresholdWidth) - 1
  int out_mul;           // 1 << (log2(fftSize) + 1): position of the threshold field in the output beat
  int thr_sext;          // 32 - thresholdWidth: sign extension of the threshold reported in a detection
  Detection *det_list;
  uint32_t *det_count;
  uint32_t det_cap;
  int skew_ns;           // A/B knob (RSP_FAST_SKEW_NS): CTAs of the second half of the grid start this much later, so that the CTAs
                         // sharing an SM are in different phases of their rows from the start
  // PLFG -> NCO fused into the first pass (rsp_run_from_generator as ONE launch, no beats in HBM): sample k of the generator's stream has
  // the NCO phase (gen_phase[k mod gen_period] + (k div gen_period) * gen_pp) mod 2^phaseWidth (rsp_host.hpp, GenPlan::phase_tab) and
  // its beat is gen_full[phase] (Tables::nco_full, built with nco_lookup); row r of the call starts at sample gen_k0 + r * N
  const uint16_t *gen_phase;
  const uint32_t *gen_full;   // null: beats come from `in`
  uint32_t gen_period, gen_pp, gen_mask;
  unsigned long long gen_k0;
  int l2_prefetch;       // 1: one thread asks the L2 for the CTA's next row (cp.async.bulk.prefetch.L2) while this row is transformed
  int stage;             // 1: the next row's beats are fetched with cp.async into a staging area behind the CTA's rows while this
                         // row's CFAR phase runs (chain_fixed_fast_kernel); needs 16-byte aligned rows and N more words per chirp
  // binPoint >= 2 (TW3 kernels; FftMagCfarChainVanillaApp has 12): the 3-multiplier twiddle product with per-product trim of MODE_FAST3
  // (rsp_device.cuh, cmul) on the raw twiddles, and the CFAR threshold with the operands' binary points (cfar_bin)
  const int2 *tw3;       // raw (cos, -sin) in Q2.14, same order and offsets as tw
  int s1, s2, h1, h1m, h2, h2m;   // product trim: >> s1 with h1 = 1 << (s1 - 1) (h1m = h1 - 1 for negated operands), result trim >> s2 with h2 / h2m
  // BPC kernels (the CFAR operands carry binary points, cfar_bin): linear thr = (t * scaler) >> thr_sh, log thr = t * t_mul + sc_pre, both at binary
  // point tbp; peak = thr - cut * 2^(tbp - inBP) < 0 (ncut_mul = -2^(tbp - inBP)); threshold field = thr >> out_sh (tbp -> thresholdBP)
  int thr_sh, ncut_mul, out_sh, t_mul, sc_pre, bpc;
  int cfar_only;  // OS / GOS bit-plane kernels: the rows are magnitude profiles (AXI4CFARBlock alone, rsp_cfar_run): no transform, the cells are the beats' low 16 bits
  int pg_full;   // peak grouping (PG kernels): 1 = guardCells equals the generator's guardWindowSize, the left neighbour is read (CFARCoreWithMem.scala:182-184)
};

// Exchange-buffer index of point p.  The pads are chosen per size by the shared-memory model in tests/test_kernel_layout.py so
// that every pass of the size's decomposition (strided 32-bit stores / loads, 128-bit loads of 4 consecutive points in the
// last pass) is free of bank conflicts (N = 8192: 1.17 wavefronts per access, the best an additive pad reaches).  All pads are
// multiples of 4 words from bit 5 up, so groups of 4 points stay contiguous and 16-byte aligned, and the index is additive over
// the disjoint bit fields it is used with (thread base | point in the group | group).
template <int S> __host__ __device__ constexpr int fidx(int p) {
  return S == 11 ? p + ((p >> 5) << 2) + ((p >> 8) << 3)
       : S == 12 ? p + ((p >> 5) << 2) + ((p >> 7) << 3)
       : S == 13 ? p + ((p >> 5) << 3)
                 : p + ((p >> 6) << 2);
}
// Index (in 16-bit units) of bin n in the magnitude array: 8 bins of padding per 64 (and per 1024 at N = 16384) spread the
// bit-reversed scatter of the last pass over the banks (N = 16384: 24 instead of 144 wavefronts per warp and row) and keep a
// thread's 16 consecutive bins contiguous and 16-byte aligned; additive over tl + j * TPC.
template <int S> __host__ __device__ constexpr int mhx(int n) { return n + ((n >> 6) << 3) + (S == 14 ? ((n >> 10) << 3) : 0); }

// LP = log2 of the points a thread owns.  16 points everywhere but at N = 16384 (RSP_FAST_LP14, default 5): there 1024 threads x 16
// points is ONE CTA per SM with every warp in the same phase of the row (issue slots 59 % busy, profiles/r01_i_ncu_full_fx14_kernel.txt);
// 512 threads x 32 points (the passes stay 4 + 4 + 4 + 2 stages on groups of 16 points, two / eight groups per thread) gives two
// independent CTAs per SM at the same 64 registers per thread, each with its own barriers, so the rows of an SM drift out of phase.
#ifndef RSP_FAST_LP14
#define RSP_FAST_LP14 5
#endif
constexpr int fast_lp(int S) { return S == 14 ? RSP_FAST_LP14 : 4; }
template <int S, int LP = 4> struct FGeo {
  static constexpr int N = 1 << S;
  static constexpr int LPTS = LP;
  static constexpr int PTS = 1 << LP;
  static constexpr int TPC = N / PTS;                        // threads per chirp
#ifndef RSP_FAST_CTA
#define RSP_FAST_CTA 128
#endif
  static constexpr int CPC = TPC >= RSP_FAST_CTA ? 1 : RSP_FAST_CTA / TPC;     // chirps per CTA
  static constexpr int THREADS = TPC * CPC;
  static constexpr int XB0 = (fidx<S>(N - 1) + 4) & ~3;
  static constexpr int XB = XB0 >= N + 4 ? XB0 : N + 4;       // words of the exchange buffer (also holds the N + 1 prefix sums)
  static constexpr int XM = (mhx<S>(N - 1) / 2 + 4) & ~3;     // words of the magnitude array (int16 per bin, padded)
  static constexpr int ROW = XB + XM;                        // words per chirp
  static constexpr int NW = TPC >= 32 ? TPC / 32 : 1;        // warps per chirp (below 512 points several chirps share a warp)
  static constexpr int LPW = TPC >= 32 ? 32 : TPC;           // lanes of one chirp inside a warp
  // Design study, off by default (-DRSP_FAST_TWSMEM_S=14 turns it on): at N = 16384 the twiddle tables of the first pass (stageN = N
  // and N / 4: 96 + 24 KB) do not stay in L1 next to a 105 KB row, every row re-reads them from L2.  Keeping them in the shared
  // memory of the SM's one CTA was measured 8 % SLOWER than reading them through L1/L2 with the cp.async staging on (151.0 vs
  // 164.7 Gsamples/s, profiles/r01_l_tw_smem_ab.txt): the kernel is short of shared-memory issue slots (mio_throttle), not of L2.
#ifndef RSP_FAST_TWSMEM_S
#define RSP_FAST_TWSMEM_S 15
#endif
  static constexpr int TWE = S >= RSP_FAST_TWSMEM_S ? 3 * (N / 4) + 3 * (N / 16) : 0;   // entries kept in shared memory
  static constexpr int TWS = 2 * TWE;                        // words
};

// Sign extensions and shifts go through inline PTX on purpose: when the compiler can prove a value fits 16 bits it rewrites the
// rounded average (a + c + 1) >> 1 into the overflow-free (a | c) - ((a ^ c) >> 1) form and narrows the arithmetic, which
// costs about twice the instructions of the plain 32-bit IADD3 + SHF.
__device__ __forceinline__ int sext16(int v) { int r; asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(r) : "r"(v)); return r; }
__device__ __forceinline__ int hi16s(uint32_t v) { int r; asm("prmt.b32 %0, %1, 0, 0xbb32;" : "=r"(r) : "r"(v)); return r; }
__device__ __forceinline__ int sar(int v, int n) { int r; asm("shr.s32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n)); return r; }
__device__ __forceinline__ uint32_t pack2(int re, int im) { return __byte_perm((uint32_t)im, (uint32_t)re, 0x5410); }
// 16-byte asynchronous copy global -> shared, past L1 (LDGSTS.BYPASS.128): no registers are held while the line is in flight
__device__ __forceinline__ void cp_async16(void *dst_shared, const void *src_global) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t)__cvta_generic_to_shared(dst_shared)), "l"(src_global) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// sa * A + sc * C with compile-time signs, as ONE IMAD (FMA pipe).  The multiplier is a run-time +-1 so that ptxas keeps the
// IMAD instead of strength-reducing it to an IADD3: the ALU pipe (PRMT, SHF, compares) is the bound of this kernel and the
// FMA pipe is half empty, so every add that can move over does.  ptxas folds the neg into the IMAD's addend modifier.
__device__ __forceinline__ int fadd(int sa, int A, int sc, int C, int one, int m1) {
  int r;

  const int mul = sa > 0 ? one : m1;
  if (sc > 0) asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(A), "r"(mul), "r"(C));
  else asm("{.reg .s32 t; neg.s32 t, %3; mad.lo.s32 %0, %1, %2, t;}" : "=r"(r) : "r"(A), "r"(mul), "r"(C));
  return r;
}

// Scaled butterfly, RoundHalfUp (generators/sdf-fft/src/main/scala/SDFFFTUtil.scala:51-61, SDFChainRadix22.scala:459-464), on
// values held with compile-time signs: A = sa * a, C = sc * c.
//   -s = (-a - c) >> 1           floor of the negated sum = minus the half-up rounded average: the +1 needs no third addend
//    d = wrap16(s - c)           because (a - c + 1) >> 1 == ((a + c + 1) >> 1) - c
// The sum position then holds -s (sign -1) and the difference position d (sign +1).  NEGD: hold wrap16(-d) instead, which is
// what the x(-j) rotation wants in its imaginary part (wrap16(-wrap16(d)) == wrap16(-d): d never is -32768 unwrapped).
// WRAP = false only where |d| is all that is used afterwards (last stage before the magnitude).
__device__ __forceinline__ void fbf1(int sa, int &A, int sc, int &C, bool negd, bool wrap, int one, int m1) {
  const int sn = sar(fadd(-sa, A, -sc, C, one, m1), 1);
  const int d = negd ? fadd(1, sn, sc, C, one, m1) : fadd(-1, sn, -sc, C, one, m1);
  A = sn;
  C = wrap ? sext16(d) : d;
}

// x <- x * W for a point held with sign sg (both components), W = (c, s) as stored (<< 2): 8 bytes per twiddle -- the twiddle
// loads are a third (N = 1024) to a half (N = 16384) of the L1 traffic of the kernel.  The product is rounded half up at bit 14
// and wrapped to 16 bits by taking bits 16..31 of the 32-bit wrap-around sum; the subtractions are IMADs with a negated addend.
//   sg > 0:              re = (xr c - (xi s - R)) >> 16,  im = (xr s + (xi c + R)) >> 16,  R = 0x8000
//   sg < 0, TRUE_OUT:    re = (xi s - (xr c - R)) >> 16,  im = (-(xr s + (xi c - R))) >> 16   (one more IMAD for the negation)
//   sg < 0, !TRUE_OUT:   the result stays negated, -floor((P + R) / 2^16) = floor((-P + R - 1) / 2^16): the sg > 0 formulas
//                        with R - 1.  Only where nothing can wrap (quiet rows: -(-32768) has no 16-bit form) and the sign
//                        can be carried on (not at the end of a pass: the exchange buffer holds true values).
__device__ __forceinline__ int imad(int x, int y, int z) { int r; asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(y), "r"(z)); return r; }
__device__ __forceinline__ int imad_sub(int x, int y, int z) {   // x * y - z
  int r; asm("{.reg .s32 t; neg.s32 t, %3; mad.lo.s32 %0, %1, %2, t;}" : "=r"(r) : "r"(x), "r"(y), "r"(z)); return r;
}
#ifndef RSP_FAST_TW_VOLATILE
#define RSP_FAST_TW_VOLATILE 0   // A/B knob: 1 = every group of a pass re-reads its twiddles (no values live across groups)
#endif
__device__ __forceinline__ int2 ldg_tw_volatile(const int2 *p) {
  int2 w; asm volatile("ld.global.nc.v2.s32 {%0, %1}, [%2];" : "=r"(w.x), "=r"(w.y) : "l"(p)); return w;
}
// rnd = {0x8000, -0x8000, 0x7FFF, -0x7FFF} come from the kernel arguments so that they are constant-bank operands of the IMADs
// (written as literals the compiler splits them off into VIADDs on the ALU pipe, the pipe that bounds the kernel)
// RAW: the products are left unshifted (the 16-bit results sit in bits 16..31 of xr / xi): the caller packs the high halves straight
// into the exchange word (one PRMT instead of two shifts and a PRMT)
template <bool TRUE_OUT, bool RAW = false>
__device__ __forceinline__ void fcmul(int &sg, int &xr, int &xi, const int2 w, int m1, const int (&rnd)[4]) {
  const int c = w.x, s = w.y;
  int r, i;
  if (sg > 0 || !TRUE_OUT) {
    const int t = imad(xi, s, sg > 0 ? rnd[1] : rnd[3]), u = imad(xi, c, sg > 0 ? rnd[0] : rnd[2]);
    r = imad_sub(xr, c, t); i = imad(xr, s, u);
  } else {
    const int t = imad(xr, c, rnd[1]), u = imad(xi, c, rnd[1]);
    r = imad_sub(xi, s, t);
    i = imad(imad(xr, s, u), m1, 0);
    sg = 1;
  }
  if (RAW) { xr = r; xi = i; } else { xr = sar(r, 16); xi = sar(i, 16); }
}

// The same scaled butterfly producing both outputs with their TRUE sign (used by the last stage of a pass, whose outputs leave the
// registers): s = (a + c + 1) >> 1 with the +1 as the third addend of one IADD3 (ALU pipe), d = s - c.  Same three instructions per
// component as fbf1, and nothing is left to un-negate before the exchange store.
#ifndef RSP_FAST_TRUE_LAST
#define RSP_FAST_TRUE_LAST 1   // A/B knob: 0 = the negated form everywhere + un-negating IMADs at the end of a pass (round 1)
#endif
__device__ __forceinline__ void fbf_true(int sa, int &A, int sc, int &C, bool wrap, int one, int m1) {
  const int t = (sa > 0 ? A : -A) + (sc > 0 ? C : -C) + 1;
  const int sv = sar(t, 1);
  const int d = fadd(1, sv, -sc, C, one, m1);
  A = sv;
  C = wrap ? sext16(d) : d;
}

// The twiddle product of the chains with binPoint >= 2 (MODE_FAST3: DspContext 3-multiplier form, every product trimmed to the context's
// binary point, the result trimmed to the data's, both RoundHalfUp; rsp_device.cuh cmul):
//   k1 = (a (br + bi) + h1) >> s1, k2 = ((a + b) bi + h1) >> s1, k3 = ((b - a) br + h1) >> s1, re = (k1 - k2 + h2) >> s2, im = (k1 + k3 + h2) >> s2
// For a point held negated the same formulas with h - 1 give the negated result exactly: -((P + h) >> s) == (-P + h - 1) >> s.
template <bool TRUE_OUT>
__device__ __forceinline__ void fcmul3(const FastArgs &a, int &sg, int &xr, int &xi, const int2 w) {
  const int br = w.x, bi = w.y, bs = br + bi;
  int A = xr, B = xi, hh1 = a.h1, hh2 = a.h2;
  if (sg < 0) {
    if (TRUE_OUT) { A = -A; B = -B; sg = 1; }
    else { hh1 = a.h1m; hh2 = a.h2m; }
  }
  const int k1 = sar(imad(A, bs, hh1), a.s1), k2 = sar(imad(A + B, bi, hh1), a.s1), k3 = sar(imad(B - A, br, hh1), a.s1);
  xr = sar(k1 - k2 + hh2, a.s2); xi = sar(k1 + k3 + hh2, a.s2);
}

// JPL magnitude approximation (generators/logMagMux/src/main/scala/LogMagMux.scala:92-110), stored to 16 bits by the caller
__device__ __forceinline__ int fjpl(int re, int im) {
  const int ar = abs(re), ai = abs(im);
  const int u = max(ar, ai), v = min(ar, ai);
  return max(u + (v >> 3), (u - (u >> 3)) + (v >> 1));
}

// One pass = m radix-2 stages in registers on 2^m points Dl apart (same decomposition as rsp_device.cuh, fx_pass_*); a thread
// owns G = PTS >> m such groups per pass.  Thread tl of the chirp works on the groups vt = g * TPC + tl, so a warp touches
// consecutive words in every pass.  fast_group is ONE group (g is a compile-time value after unrolling).
// WRAP = false: the row was found "quiet" (see quiet_beats): no value of the transform can reach +-32768, so the 16-bit wrap
// of the butterfly differences is the identity and its PRMT is left out (13% of the ALU-pipe instructions of the kernel).
// FFTO (the FFT block alone, fft_fixed_fast_kernel): the last pass ends like the others -- true signs -- and leaves the packed output
// beats Cat(re16, im16) in `mags` instead of the magnitudes.
template <int S, int LP, int K, bool WRAP, bool TWSM = false, bool FFTO = false, bool TW3 = false>
__device__ __forceinline__ void fast_group(const FastArgs &a, int tl, int g, const uint32_t (&pre)[16], uint32_t *xb, int (&mags)[1 << LP],
                                           const int2 *tws = nullptr) {
  using Gm = FGeo<S, LP>;
  constexpr int m = fx_pass_size(S, K), i0 = fx_pass_start(S, K), LDl = S - i0 - m, Dl = 1 << LDl;
  constexpr int R = 1 << m, TPC = Gm::TPC;
  constexpr bool first = (K == 0), last = (K == fx_n_passes(S) - 1);
  static_assert((i0 & 1) == 0, "passes start on even stages");
  static_assert(!first || m == 4, "the first pass takes 16 beats per group");
  // exchange-buffer address of point (g, rr) = xf[fidx(group offset) + fidx(rr * Dl)]: one run-time base per pass, the rest
  // folds into the instruction's immediate (fidx is additive over these disjoint bit fields).  Dl <= TPC: the groups of a thread
  // are TPC * R points apart and share low = tl mod Dl; Dl > TPC (first pass with 32 points per thread): they are TPC apart and
  // low = g * TPC + tl.
  constexpr bool wide = Dl > TPC;
  const int low0 = tl & (Dl - 1);
  uint32_t *xf = xb + fidx<S>(((tl >> LDl) << (m + LDl)) | low0);
  const int go = fidx<S>(wide ? g * TPC : g * TPC * R);
  const int klow = wide ? g * TPC : 0;     // compile-time part of low
  int re[R], im[R];
  if (last && !first) {   // R consecutive points: 128-bit loads
#pragma unroll
    for (int q = 0; q < R / 4; ++q) {
      const uint4 v = *reinterpret_cast<const uint4 *>(&xf[go + fidx<S>(4 * q)]);
      const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { re[4 * q + k] = hi16s(wd[k]); im[4 * q + k] = sext16((int)wd[k]); }
    }
  } else {
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const uint32_t wd = first ? pre[rr] : xf[go + fidx<S>(rr * Dl)];   // first pass: the group's 16 beats, p = tl + (rr * G + g) * TPC
      re[rr] = hi16s(wd); im[rr] = sext16((int)wd);
    }
  }
  int sg[R];   // sign each point is held with (compile-time after unrolling): value = sg * register
  bool raw[R]; // the point's components sit in bits 16..31 of its registers (unshifted twiddle product of the pass's last stage)
#pragma unroll
  for (int rr = 0; rr < R; ++rr) { sg[rr] = 1; raw[rr] = false; }
#pragma unroll
  for (int s = 0; s < m; ++s) {
    const int i = i0 + s;
    const int dist = 1 << (m - 1 - s);
    const bool rot = (i != S - 1) && ((i & 1) == 0);   // x(-j) on the last quarter of each 4D block follows (SDFChainRadix22.scala:319-337)
    const bool wrap = WRAP && !(last && s == m - 1);    // the magnitude only takes |d| (|wrap16(32768)| == 32768)
    const bool true_stage = RSP_FAST_TRUE_LAST && (!last || FFTO) && s == m - 1;   // the outputs leave the registers: true signs, raw products
    static_assert(last || (((i0 + m - 1) & 1) == 1), "a pass that is not the last ends on an odd stage (no rotation pending)");
#pragma unroll
    for (int rr = 0; rr < R; ++rr)
      if (!(rr & dist)) {
        const int rc = rr + dist;
        if (true_stage) {
          fbf_true(sg[rr], re[rr], sg[rc], re[rc], wrap, a.one, a.m1);
          fbf_true(sg[rr], im[rr], sg[rc], im[rc], wrap, a.one, a.m1);
          sg[rr] = 1; sg[rc] = 1;
          continue;
        }
        const bool mj = rot && (rc & (dist >> 1));      // the difference point is rotated: (re, im) <- (im, -re)
        fbf1(sg[rr], re[rr], sg[rc], re[rc], mj, wrap, a.one, a.m1);
        fbf1(sg[rr], im[rr], sg[rc], im[rc], false, wrap, a.one, a.m1);
        if (mj) { const int t = re[rc]; re[rc] = im[rc]; im[rc] = t; }
        sg[rr] = -1; sg[rc] = 1;
      }
    if (i == S - 1 || rot) continue;
    const int L = S - i + 1;
    const int q = 1 << (L - 2);
    constexpr bool smt = TWSM && first;                      // this pass's tables sit in shared memory: [stageN = N | N / 4]
    const int2 *tab = smt ? tws + (i == 1 ? 0 : 3 * (Gm::N / 4)) : (TW3 ? a.tw3 : a.tw) + a.tw_off[L];
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int quarter = ((rr >> (m - 1 - s)) & 1) + 2 * ((rr >> (m - s)) & 1);
      if (quarter) {
        const int k = (rr & (dist - 1)) * Dl + klow + low0;   // p mod q, q = stageN / 4 = Dl * dist
        const int slot = quarter == 1 ? 0 : (quarter == 2 ? 1 : 2);
        // several groups per pass (32 points per thread) use the same twiddles: left to CSE, the compiler keeps all 15 pairs of a
        // pass live across the groups and spills the row loop's invariants; a volatile load per group re-reads them from L1 instead
        const int2 w = smt ? tab[slot * q + k] : ((RSP_FAST_TW_VOLATILE && LP > 4 && !first) ? ldg_tw_volatile(&tab[slot * q + k]) : __ldg(&tab[slot * q + k]));
        if (TW3) {   // the result is a plain value: in loud rows it is wrapped like every other 16-bit assignment
          if (true_stage || WRAP || (s == m - 1 && !last)) fcmul3<true>(a, sg[rr], re[rr], im[rr], w); else fcmul3<false>(a, sg[rr], re[rr], im[rr], w);
          if (WRAP) { re[rr] = sext16(re[rr]); im[rr] = sext16(im[rr]); }
        }
        else if (true_stage) { fcmul<true, true>(sg[rr], re[rr], im[rr], w, a.m1, a.rnd); raw[rr] = true; }
        else if (WRAP || (s == m - 1 && !last)) fcmul<true>(sg[rr], re[rr], im[rr], w, a.m1, a.rnd);
        else fcmul<false>(sg[rr], re[rr], im[rr], w, a.m1, a.rnd);
      }
    }
  }
  if (!last || FFTO) {   // the exchange buffer (the output) holds true values: un-negate what is still held negated
#pragma unroll
    for (int rr = 0; rr < R; ++rr)
      if (sg[rr] < 0) { re[rr] = fadd(-1, re[rr], 1, 0, a.one, a.m1); im[rr] = fadd(-1, im[rr], 1, 0, a.one, a.m1); }
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const uint32_t wd = raw[rr] ? __byte_perm((uint32_t)im[rr], (uint32_t)re[rr], 0x7632) : pack2(re[rr], im[rr]);
      if (last) mags[g * R + rr] = (int)wd; else xf[go + fidx<S>(rr * Dl)] = wd;
    }
  } else {
#pragma unroll
    for (int rr = 0; rr < R; ++rr) mags[g * R + rr] = fjpl(re[rr], im[rr]);
  }
}

// Bit reversal = scatter of the last pass's magnitudes (16 bits each) into the magnitude array.
// bin of point p = (vt << m) + rr, vt = g * TPC + tl, is bitrev_S(p) = rev_m(rr) << (S - m) | rev_{S-LP}(tl) << (LP - m) | rev_{LP-m}(g):
// for a fixed rr the G groups of the thread are G consecutive bins -> one G * 16-bit store per rr
template <int S, int LP>
__device__ __forceinline__ void fast_scatter(int tl, const int (&mags)[1 << LP], uint32_t *xm) {
  constexpr int K = fx_n_passes(S) - 1, m = fx_pass_size(S, K), R = 1 << m, G = (1 << LP) >> m;
  const uint32_t rtl = (__brev((uint32_t)tl) >> (32 - (S - LP))) << (LP - m);
  uint16_t *xm16 = reinterpret_cast<uint16_t *>(xm) + mhx<S>((int)rtl);
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    const uint32_t rrr = (__brev((uint32_t)rr) >> (32 - m)) << (S - m);   // compile-time
    const int bin0 = mhx<S>((int)rrr);                                     // mhx is additive over rrr | rtl
    auto pk = [&](int g0, int g1) { return __byte_perm((uint32_t)mags[g0 * R + rr], (uint32_t)mags[g1 * R + rr], 0x5410); };
    if (G == 8) {          // groups in bin order: rev3(g) = 0..7 <- g = 0,4,2,6,1,5,3,7
      *reinterpret_cast<uint4 *>(&xm16[bin0]) = make_uint4(pk(0, 4), pk(2, 6), pk(1, 5), pk(3, 7));
    } else if (G == 4) {   // rev2(g) = 0,1,2,3 <- g = 0,2,1,3
      *reinterpret_cast<uint2 *>(&xm16[bin0]) = make_uint2(pk(0, 2), pk(1, 3));
    } else if (G == 2) {
      *reinterpret_cast<uint32_t *>(&xm16[bin0]) = pk(0, 1);
    } else {
      static_assert(G == 8 || G == 4 || G == 2 || G == 1, "groups of the last pass");
      xm16[bin0] = (uint16_t)mags[rr];
    }
  }
}

// a whole pass whose groups need no beats from global memory (every pass but the first, and the first one with 16 points per thread)
template <int S, int LP, int K, bool WRAP, bool TWSM = false, bool FFTO = false, bool TW3 = false>
__device__ __forceinline__ void fast_pass(const FastArgs &a, int tl, const uint32_t (&pre)[16], uint32_t *xb, uint32_t *xm, const int2 *tws = nullptr,
                                          int *wout = nullptr) {
  constexpr int m = fx_pass_size(S, K), G = (1 << LP) >> m;
  constexpr bool last = (K == fx_n_passes(S) - 1);
  static_assert(K > 0 || G == 1, "first pass with several groups: fast_first_pass_groups");
  int mags[1 << LP];
  // last pass: the groups in the order in which their magnitudes pair up in the scatter's packed words (rev(g) ascending), so that
  // a pair is packed as soon as it exists instead of 2^LP magnitudes staying live
#pragma unroll
  for (int gi = 0; gi < G; ++gi) {
    const int g = !last ? gi : (G == 8 ? ((gi & 1) << 2 | (gi & 2) | (gi >> 2)) : (G == 4 ? ((gi & 1) << 1 | (gi >> 1)) : gi));
    fast_group<S, LP, K, WRAP, TWSM, FFTO, TW3>(a, tl, g, pre, xb, mags, tws);
    // keep the next group's shared-memory loads behind this group's stores: hoisted, 32 points are live and the 64-register
    // budget spills (the other warps of the SM hide the load latency, not this thread's next group)
    // (ptxas moves shared-memory accesses with distinct immediate offsets across a compiler-level fence; WARPSYNC holds them)
    if (!last && gi + 1 < G) __syncwarp();
  }
  if (last) {
    if (FFTO) {   // the output beats stay in registers: the exchange buffer they go to is still being read by the other threads' last pass
#pragma unroll
      for (int e = 0; e < (1 << LP); ++e) wout[e] = mags[e];
    } else fast_scatter<S, LP>(tl, mags, xm);
  }
}

// Beats j = 0 .. 15 of group g of the thread, p = tl + (j * G + g) * TPC (warp-contiguous 128-byte lines), times the window when
// there is one: bits 16..31 of beat * (window << (16 - bp)) are wrap16(floor(beat * window >> bp)), the truncating assignment of
// SDFFFT.scala:233-234; the windowed beats are packed again so that everything behind (the quiet test too) sees plain beats.
template <int S, int LP>
__device__ __forceinline__ void window_beats(const FastArgs &a, int tl, int g, uint32_t (&pre)[16]) {
  constexpr int TPC = FGeo<S, LP>::TPC, G = (1 << LP) / 16;
  if (a.win4 != nullptr) {
    const int32_t *gw = a.win4 + tl + g * TPC;
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) {
      const int w = __ldg(&gw[rr * G * TPC]);
      pre[rr] = __byte_perm((uint32_t)(sext16((int)pre[rr]) * w), (uint32_t)(hi16s(pre[rr]) * w), 0x7632);
    }
  }
}
// position of a sample in the generator's periodic schedule: k = per * gen_period + q
struct GenRow { uint32_t q, per; };
__device__ __forceinline__ void gen_advance(uint32_t &q, uint32_t &per, uint32_t dq, uint32_t dper, uint32_t P) {
  q += dq; per += dper;
  if (q >= P) { q -= P; ++per; }
}
template <int S, int LP>
__device__ __forceinline__ void load_beats(const FastArgs &a, const uint32_t *grow, const GenRow &gr, int tl, int g, uint32_t (&pre)[16]) {
  constexpr int TPC = FGeo<S, LP>::TPC, G = (1 << LP) / 16;
  if (a.gen_full != nullptr) {   // the beats of this group from the PLFG -> NCO schedule: points p = tl + (rr * G + g) * TPC of the row
    const uint32_t P = a.gen_period;
    const uint32_t dq1 = (uint32_t)TPC % P, dper1 = (uint32_t)TPC / P, dqG = (uint32_t)(G * TPC) % P, dperG = (uint32_t)(G * TPC) / P;
    uint32_t q = gr.q, per = gr.per;
#pragma unroll
    for (int i = 0; i < g; ++i) gen_advance(q, per, dq1, dper1, P);
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) {
      const uint32_t ph = ((uint32_t)__ldg(&a.gen_phase[q]) + per * a.gen_pp) & a.gen_mask;
      pre[rr] = __ldg(&a.gen_full[ph]);
      gen_advance(q, per, dqG, dperG, P);
    }
  } else {
    const uint32_t *gin = grow + tl + g * TPC;
#pragma unroll
    for (int rr = 0; rr < 16; ++rr) pre[rr] = __ldg(&gin[rr * G * TPC]);
  }
  window_beats<S, LP>(a, tl, g, pre);
}

// Barrier between the phases of a row.  The chirps of a CTA share nothing, so a named barrier per chirp (bar.sync c + 1, TPC)
// would do -- measured on B200 it is 7% SLOWER than the plain CTA barrier (204 vs 219 Gsamples/s), so the CTA barrier stays.
template <int S> __device__ __forceinline__ void chirp_sync(int) { __syncthreads(); }

// A row is "quiet" when every input component is within +-QUIET_MAX.  Then |x| <= 32739 for every complex input, a scaled
// butterfly raises the largest modulus by at most sqrt(2)/2 (the two roundings), a twiddle product by at most
// 32767 * 2^-14 * sqrt(2)/2 + sqrt(2)/2 (twiddle quantisation + rounding), the x(-j) keeps it: after S <= 14 stages and
// 6 products no modulus, hence no component, exceeds 32739 + 14 * 0.71 + 6 * 2.2 < 32767, and nothing ever wraps.
// Rows that are not quiet (full-scale inputs, the -32768 corners) take the passes with the wraps; both give the same bits.
#define RSP_QUIET_MAX 23100
__device__ __forceinline__ bool quiet_beats(const uint32_t (&pre)[16]) {
  uint32_t mx = pre[0], mn = pre[0];
#pragma unroll
  for (int k = 1; k < 15; k += 2) { mx = __vimax3_s16x2(mx, pre[k], pre[k + 1]); mn = __vimin3_s16x2(mn, pre[k], pre[k + 1]); }
  mx = __vmaxs2(mx, pre[15]); mn = __vmins2(mn, pre[15]);
  const uint32_t lim = (uint32_t)RSP_QUIET_MAX * 0x10001u, nlim = (uint32_t)(0x10000 - RSP_QUIET_MAX) * 0x10001u;
  return __vmins2(mx, lim) == mx && __vmaxs2(mn, nlim) == mn;
}

// The first pass works on the thread's own beats, so each warp decides for itself (per group of 16 beats); the later passes mix
// the whole chirp and decide by the chirp's flag, which every warp of the chirp has or-ed into before the barrier that ends the
// first pass.
template <int S, int LP, int K, bool FFTO = false, bool TW3 = false>
__device__ __forceinline__ void fast_passes_from(const FastArgs &a, int c, int tl, bool quiet, const uint32_t (&pre)[16], uint32_t *xb, uint32_t *xm,
                                                 int *wout = nullptr) {
  if (quiet) fast_pass<S, LP, K, false, false, FFTO, TW3>(a, tl, pre, xb, xm, nullptr, wout);
  else fast_pass<S, LP, K, true, false, FFTO, TW3>(a, tl, pre, xb, xm, nullptr, wout);
  chirp_sync<S>(c);
  if constexpr (K + 1 < fx_n_passes(S)) fast_passes_from<S, LP, K + 1, FFTO, TW3>(a, c, tl, quiet, pre, xb, xm, wout);
}

// pre: the thread's 16 beats (LP = 4; loaded or staged by the caller).  grow: the row in global memory (LP = 5: each of the two
// groups of the first pass loads its own 16 beats, so that no more than one group's beats are live at a time).
template <int S, int LP, bool TWSM = false, bool FFTO = false, bool TW3 = false>
__device__ __forceinline__ void fast_passes(const FastArgs &a, int c, int tl, uint32_t (&pre)[16], const uint32_t *grow, const GenRow &gr, uint32_t *xb,
                                            uint32_t *xm, int *loud, const int2 *tws = nullptr, int *wout = nullptr) {
  static_assert(fx_pass_size(S, 0) == 4, "the first pass works on groups of 16 beats");
  constexpr int G0 = (1 << LP) / 16;
  if constexpr (G0 == 1) {
    const bool warp_quiet = __all_sync(0xffffffffu, quiet_beats(pre));
    if (!warp_quiet && (tl & (FGeo<S, LP>::LPW - 1)) == 0) loud[c] = 1;   // (a warp that holds several chirps flags them all)
    if (warp_quiet) fast_pass<S, LP, 0, false, TWSM, false, TW3>(a, tl, pre, xb, xm, tws); else fast_pass<S, LP, 0, true, TWSM, false, TW3>(a, tl, pre, xb, xm, tws);
  } else {
    int mags[1 << LP];   // not used by a first pass
    bool all_quiet = true;
#pragma unroll
    for (int g = 0; g < G0; ++g) {
      load_beats<S, LP>(a, grow, gr, tl, g, pre);
      const bool warp_quiet = __all_sync(0xffffffffu, quiet_beats(pre));
      all_quiet = all_quiet && warp_quiet;
      if (warp_quiet) fast_group<S, LP, 0, false, TWSM, false, TW3>(a, tl, g, pre, xb, mags, tws); else fast_group<S, LP, 0, true, TWSM, false, TW3>(a, tl, g, pre, xb, mags, tws);
    }
    if (!all_quiet && (tl & (FGeo<S, LP>::LPW - 1)) == 0) loud[c] = 1;
  }
  chirp_sync<S>(c);
  if constexpr (fx_n_passes(S) > 1) {
    const bool quiet = loud[c] == 0;
    fast_passes_from<S, LP, 1, FFTO, TW3>(a, c, tl, quiet, pre, xb, xm, wout);     // ends with a barrier: every thread has read the flag
    if (tl == 0) loud[c] = 0;                                 // next written before the first barrier of the next row
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// CFAR (CFARCoreWithMem.scala:102-189).  P[n] = sum of x[0..n-1] lives in the exchange buffer (dead after the last FFT pass);
// lagging sum of bin n = P[n+G+W+1] - P[n+G+1], leading sum = P[n-G] - P[n-G-W].

// exclusive prefix of `total` over the TPC threads of one chirp: warp shuffles, then one shared-memory hop across the warps
template <int S, int LP>
__device__ __forceinline__ int chirp_scan_exclusive(int total, int c, int tl, int *wsum) {
  constexpr int NW = FGeo<S, LP>::NW, LPW = FGeo<S, LP>::LPW;
  const int lane = tl & (LPW - 1);
  int incl = total;
#pragma unroll
  for (int d = 1; d < LPW; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d, LPW);   // segments of LPW lanes: the chirps that share a warp do not mix
    if (lane >= d) incl += v;
  }
  int offset = 0;
  if (NW > 1) {
    const int w = tl >> 5;
    if (lane == 31) wsum[c * NW + w] = incl;
    chirp_sync<S>(c);
#pragma unroll
    for (int h = 0; h < NW - 1; ++h) offset += (h < w) ? wsum[c * NW + h] : 0;
  }
  return offset + incl - total;
}

// Threshold, peak and output beat of bin n = tl + j * TPC; returns the peak bit.  pA..pD point at P[tl + offset] for the four
// window borders, so every load is [register + immediate].  EDGE: the bin may miss one of its windows (one-sided threshold,
// CFARCoreWithMem.scala:154-171); the indices are clamped so that no load leaves the chirp's arrays.
// !has_out (the caller wants the detection list only): no output beat is built or stored.
// PG: peak grouping -- a peak must also exceed its two "neighbours" as the core's guard registers deliver them: laggGuard.last = x[n + 1]
// (0 unless guardCells == guardWindowSize) and laggGuard.head = x[n + guardCells] (CFARCoreWithMem.scala:182-184; both from the lagging side).
// BPV: the operands carry a binary point (inBP = scalerBP = thresholdBP = b >= 1): the linear threshold is the product trimmed to the
// context's binary point tbp (Floor: an arithmetic shift), the cell under test is compared at tbp, the field is the threshold at b.
template <int S, int LP, int CMODE, bool LINEAR, bool EDGE, bool PG = false, bool BPV = false>   // BPV here = the kernel's BPC
__device__ __forceinline__ int fast_cfar_bin(const FastArgs &a, int j, int tl, int n2, const int *P, const int *pA, const int *pB,
                                             const int *pC, const int *pD, const short *X, uint32_t *gout, bool has_out, const short *X0 = nullptr) {
  constexpr int N = 1 << S, TPC = FGeo<S, LP>::TPC;
  const int n = tl + j * TPC;
  int L, R;
  bool lagg_full = true, lead_full = true;
  if (EDGE) {
    const int W = a.W, G = a.G;
    lagg_full = (n + G + W <= N - 1); lead_full = (n - G - W >= 0);
    const int hi = min(n + G + W + 1, N), lo = max(n - G - W, 0);
    L = fadd(1, P[hi], -1, P[hi - W], a.one, a.m1);
    R = fadd(1, P[lo + W], -1, P[lo], a.one, a.m1);
  } else {
    L = fadd(1, pA[j * TPC], -1, pB[j * TPC], a.one, a.m1);
    R = fadd(1, pC[j * TPC], -1, pD[j * TPC], a.one, a.m1);
  }
  L = sar(L, a.div_sum); R = sar(R, a.div_sum);
  int t = CMODE == 0 ? sar(fadd(1, R, 1, L, a.one, a.m1), 1) : (CMODE == 1 ? max(L, R) : min(L, R));
  if (EDGE) t = lead_full ? (lagg_full ? t : R) : L;
  const int thr = LINEAR ? ((BPV) ? sar(t * a.scaler, a.thr_sh) : t * a.scaler) : (BPV ? imad(t, a.t_mul, a.sc_pre) : fadd(1, t, 1, a.scaler, a.one, a.m1));
  const int cut = X[mhx<S>(j * TPC)];   // X points at the thread's first bin (mhx is additive over tl + j * TPC)
  // peak = cut > thr as the sign bit of thr - cut (|thr| < 2^31 - 2^16 is checked by the host)
  int pk = BPV ? (int)((uint32_t)imad(cut, a.ncut_mul, thr) >> 31) : (int)((uint32_t)fadd(1, thr, -1, cut, a.one, a.m1) >> 31);
  if (PG) {
    const int n1 = n + 1, ng = n + a.G;
    int left = 0, right = 0;
    if (a.pg_full && (!EDGE || n1 < N)) left = X0[mhx<S>(n1)];
    if (!EDGE || ng < N) right = X0[mhx<S>(ng)];
    pk = (pk != 0 && cut > left && cut > right) ? 1 : 0;
  }
  if (has_out) {
    const int tf = BPV ? sar(thr, a.out_sh) : thr;
    const int w0 = (int)(((uint32_t)tf & a.thr_mask) * (uint32_t)a.out_mul + (uint32_t)n2);   // threshold field | bin field (n << 1)
    gout[n] = (uint32_t)fadd(1, pk, 1, w0, a.one, a.m1);
  }
  return pk;
}

// detections of one thread: bit (PTS - 1 - j) of `mask` = peak of bin tl + j * TPC.  Rare path, recomputes from shared memory.
struct FastDetArgs { int W, G, div_sum, scaler, thr_sext;   /* div_sum as in FastArgs */ uint32_t chirp, det_cap; Detection *det_list; uint32_t *det_count; int thr_sh, out_sh, t_mul, sc_pre; };
template <int S, int LP, int CMODE, bool LINEAR>
__device__ __noinline__ void fast_detections(const FastDetArgs a, int tl, uint32_t mask, const int *P, const short *X) {
  constexpr int N = 1 << S, TPC = FGeo<S, LP>::TPC, PTS = 1 << LP;
  for (int j = 0; j < PTS; ++j) {
    if (!((mask >> (PTS - 1 - j)) & 1u)) continue;
    const int n = tl + j * TPC, W = a.W, G = a.G;
    const bool lagg_full = (n + G + W <= N - 1), lead_full = (n - G - W >= 0);
    const int hi = min(n + G + W + 1, N), lo = max(n - G - W, 0);
    const int L = (int)((uint32_t)P[hi] - (uint32_t)P[hi - W]) >> a.div_sum, R = (int)((uint32_t)P[lo + W] - (uint32_t)P[lo]) >> a.div_sum;
    int t = CMODE == 0 ? ((R + L) >> 1) : (CMODE == 1 ? max(L, R) : min(L, R));
    t = lead_full ? (lagg_full ? t : R) : L;
    const int thr = (LINEAR ? (t * a.scaler) >> a.thr_sh : t * a.t_mul + a.sc_pre) >> a.out_sh;   // (without binary points: shifts 0, t_mul 1, sc_pre = scaler)
    const uint32_t slot = atomicAdd(a.det_count, 1u);
    if (slot < a.det_cap) {
      Detection d; d.chirp = a.chirp; d.bin = (uint32_t)n; d.threshold = (int)((uint32_t)thr << a.thr_sext) >> a.thr_sext; d.cut = X[mhx<S>(n)];
      a.det_list[slot] = d;
    }
  }
}

// EJ = number of leading / trailing iterations j whose bins may miss a window (ceil((W + G) / TPC), checked by the host)
// resident CTAs per SM the register budget is cut for: 1024 threads at 64 registers by default; RSP_FAST_MINB<S> (A/B knob) asks for
// fewer CTAs with more registers each
#ifndef RSP_FAST_MINB14
#define RSP_FAST_MINB14 0
#endif
#ifndef RSP_FAST_MINB10
#define RSP_FAST_MINB10 0
#endif
template <int S, int LP> constexpr int fast_min_blocks() {
  return (S == 14 && RSP_FAST_MINB14 > 0) ? RSP_FAST_MINB14 : ((S == 10 && RSP_FAST_MINB10 > 0) ? RSP_FAST_MINB10 : 1024 / FGeo<S, LP>::THREADS);
}
template <int S, int LP, int CMODE, bool LINEAR, int EJ, bool PG = false, bool TW3 = false, bool BPC = false>
__global__ void __launch_bounds__(FGeo<S, LP>::THREADS, fast_min_blocks<S, LP>()) chain_fixed_fast_kernel(const FastArgs a) {
  using Gm = FGeo<S, LP>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS;
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ int wsum[CPC * Gm::NW];
  __shared__ int loud[CPC];   // per chirp: some warp saw a beat that is not quiet (fast_passes)
  if (threadIdx.x < CPC) loud[threadIdx.x] = 0;
  __syncthreads();
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *xb = smem + (size_t)c * Gm::ROW, *xm = xb + Gm::XB;
  int *P = reinterpret_cast<int *>(xb);
  const short *X = reinterpret_cast<const short *>(xm), *Xt = X + mhx<S>(tl);
  const int *pA = P + tl + a.G + a.W + 1, *pB = P + tl + a.G + 1, *pC = P + tl - a.G, *pD = P + tl - a.G - a.W;
  const uint32_t stride = gridDim.x * CPC;
  // Staging (a.stage, 16 points per thread only): the beats of the CTA's NEXT rows are copied global -> shared by cp.async (4 x 16
  // bytes per thread, no registers in flight) once this row's transform is done, and land while its CFAR phase runs; the row loop
  // then starts from shared memory instead of waiting on HBM with every warp of the CTA at once.
  uint32_t *stg = smem + (size_t)CPC * Gm::ROW + (size_t)c * N;
  const bool stage = LP == 4 && a.stage != 0;
  constexpr bool TWSM = Gm::TWS > 0;
  const int2 *tws = reinterpret_cast<const int2 *>(smem + (size_t)CPC * Gm::ROW);   // shares the place of the staging area: one or the other
  if (TWSM) {
    int4 *dst = reinterpret_cast<int4 *>(smem + (size_t)CPC * Gm::ROW);
    const int4 *s0 = reinterpret_cast<const int4 *>(a.tw + a.tw_off[S]), *s1 = reinterpret_cast<const int4 *>(a.tw + a.tw_off[S - 2]);
    constexpr int n0 = 3 * (N / 4) / 2, n1 = 3 * (N / 16) / 2;   // int4 = two entries
    for (int e = threadIdx.x; e < n0; e += Gm::THREADS) dst[e] = __ldg(&s0[e]);
    for (int e = threadIdx.x; e < n1; e += Gm::THREADS) dst[n0 + e] = __ldg(&s1[e]);
    __syncthreads();
  }
  auto prefetch = [&](uint32_t r0) {
    const uint32_t rw = min(r0 + (uint32_t)c, a.n_rows - 1);
    const uint32_t *g = a.in + (size_t)rw * N;
#pragma unroll
    for (int k = 0; k < 4; ++k) { const int e = (k * TPC + tl) * 4; cp_async16(stg + e, g + e); }
    cp_async_commit();
  };
  if (stage) { prefetch(blockIdx.x * CPC); cp_async_wait_all(); __syncthreads(); }
  // generator fused in: schedule position of point tl of the CTA's current row, advanced by stride * N samples per iteration
  GenRow gr = {0u, 0u};
  uint32_t g_dq = 0, g_dper = 0;
  const bool gen = a.gen_full != nullptr;
  if (gen) {
    const unsigned long long k = a.gen_k0 + ((unsigned long long)(blockIdx.x * CPC + c) << S) + (unsigned)tl;
    gr.per = (uint32_t)(k / a.gen_period); gr.q = (uint32_t)(k - (unsigned long long)gr.per * a.gen_period);
    const uint32_t dk = stride << S;
    g_dper = dk / a.gen_period; g_dq = dk - g_dper * a.gen_period;
  }
  if (a.skew_ns > 0 && 2 * blockIdx.x >= gridDim.x) {
    for (int left = a.skew_ns; left > 0; left -= 1000) __nanosleep(1000);
  }
  // the loop bound is the same for every thread of the CTA (the barriers are CTA-wide); a chirp slot past the last row
  // recomputes the last row -- it stores the same beats again and reports no detections
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const bool active = row0 + c < a.n_rows;
    const uint32_t row = active ? row0 + c : a.n_rows - 1;
    const uint32_t *grow = a.in + (size_t)row * N;
    uint32_t pre[16];
    GenRow grr = gr;
    if (gen && !active) {   // a chirp slot past the last row recomputes the last row: its own schedule position
      const unsigned long long k = a.gen_k0 + ((unsigned long long)row << S) + (unsigned)tl;
      grr.per = (uint32_t)(k / a.gen_period); grr.q = (uint32_t)(k - (unsigned long long)grr.per * a.gen_period);
    }
    if (LP == 4) {
      if (stage) {
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) pre[rr] = stg[tl + rr * TPC];
        window_beats<S, LP>(a, tl, 0, pre);
      } else load_beats<S, LP>(a, grow, grr, tl, 0, pre);
    }
    fast_passes<S, LP, TWSM, false, TW3>(a, c, tl, pre, grow, grr, xb, xm, loud, tws);
    if (gen) gen_advance(gr.q, gr.per, g_dq, g_dper, a.gen_period);
    // every thread is past the first barrier of the passes: the staging area has been read and can take the next rows
    if (stage && row0 + stride < a.n_rows) prefetch(row0 + stride);
    if (a.l2_prefetch && tl == 0 && row0 + a.l2_prefetch * stride + c < a.n_rows) {   // a coming row's 4 N bytes into L2 (l2_prefetch rows ahead): its first-pass loads then miss L1 only
      const uint32_t *nxt = a.in + (size_t)(row0 + a.l2_prefetch * stride + c) * N;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(nxt), "r"(4 * N) : "memory");
    }
    // prefix sums of the thread's PTS consecutive magnitudes, 16 at a time (32 points per thread: the second read of the magnitudes
    // is cheaper than keeping them in registers across the scan)
    const uint32_t *xq = xm + mhx<S>(PTS * tl) / 2;   // the thread's PTS consecutive bins (inside one 64-bin block of the padded array)
    auto load16 = [&](int h, int (&x)[16]) {
      const uint4 v0 = *reinterpret_cast<const uint4 *>(xq + 8 * h), v1 = *reinterpret_cast<const uint4 *>(xq + 8 * h + 4);
      const uint32_t wd[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) { x[2 * k] = sext16((int)wd[k]); x[2 * k + 1] = hi16s(wd[k]); }
    };
    int x[16];
    int total = 0;
#pragma unroll
    for (int h = 0; h < PTS / 16; ++h) {
      load16(h, x);
#pragma unroll
      for (int k = 0; k < 16; ++k) total = (int)((uint32_t)x[k] * (uint32_t)a.wmul + (uint32_t)total);   // mod 2^32 on purpose
    }
    int acc = chirp_scan_exclusive<S, LP>(total, c, tl, wsum);
#pragma unroll
    for (int h = 0; h < PTS / 16; ++h) {
      if (PTS > 16) load16(h, x);
      int pv[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) { pv[k] = acc; acc = (int)((uint32_t)x[k] * (uint32_t)a.wmul + (uint32_t)acc); }
#pragma unroll
      for (int k = 0; k < 16; k += 4) *reinterpret_cast<int4 *>(&P[tl * PTS + 16 * h + k]) = make_int4(pv[k], pv[k + 1], pv[k + 2], pv[k + 3]);
    }
    if (tl == TPC - 1) P[N] = acc;
    chirp_sync<S>(c);
    uint32_t *gout = a.out + (size_t)row * N;   // formed from the kernel argument unconditionally: a select with nullptr would make the stores generic (ST.E)
    const bool has_out = a.out != nullptr;
    uint32_t mask = 0;
    int n2 = tl << 1;
#pragma unroll
    for (int j = 0; j < PTS; ++j) {
      int pk;
      if (j < EJ || j >= PTS - EJ) pk = fast_cfar_bin<S, LP, CMODE, LINEAR, true, PG, BPC>(a, j, tl, n2, P, pA, pB, pC, pD, Xt, gout, has_out, X);
      else pk = fast_cfar_bin<S, LP, CMODE, LINEAR, false, PG, BPC>(a, j, tl, n2, P, pA, pB, pC, pD, Xt, gout, has_out, X);
      mask = mask * 2u + (uint32_t)pk;
      n2 += 2 * TPC;
    }
    if (mask != 0 && a.det_list != nullptr && active) {
      const FastDetArgs da = {a.W, a.G, a.div_sum, a.scaler, a.thr_sext, a.row_base + row, a.det_cap, a.det_list, a.det_count, BPC ? a.thr_sh : 0, BPC ? a.out_sh : 0, BPC ? a.t_mul : 1, BPC ? a.sc_pre : a.scaler};
      fast_detections<S, LP, CMODE, LINEAR>(da, tl, mask, P, X);
    }
    if (stage) cp_async_wait_all();   // this thread's copies have landed; the barrier publishes everybody's
    chirp_sync<S>(c);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// The FFT block alone (AXI4FFTBlock, rsp_fft_run) on the fast passes: same transform, the bit-reversed output beats Cat(re16, im16)
// instead of the magnitudes.  The beats are scattered into the row's shared memory (the exchange buffer is dead after the last pass; its
// words and the magnitude array behind it are one region here) at wx(bin) and leave in natural order, one warp-contiguous line per store.
// A lane's run of G = 2^(LP - m) neighbouring bins starts at rev_m(rr) << (S - m) | rev_{S-LP}(tl) << (LP - m): the lanes of a warp differ in
// rev5(lane) << (S - m - 5), all in one bank without padding.  wx pads by one store width per 2^sh words with sh chosen per size so that
// the 8 (128-bit), 16 (64-bit) or 32 (32-bit stores) lanes of a store phase fall into distinct banks (tests/test_kernel_layout.py).
template <int S> __host__ __device__ constexpr int wx(int n) {
  return S == 10 ? n + ((n >> 5) << 2) : S == 11 ? n + ((n >> 4) << 1) : S == 12 ? n + (n >> 3) : S == 13 ? n + ((n >> 6) << 1) : n + ((n >> 9) << 2);
}
template <int S, int LP>
__device__ __forceinline__ void fast_scatter_words(int tl, const int (&w)[1 << LP], uint32_t *xw) {
  constexpr int K = fx_n_passes(S) - 1, m = fx_pass_size(S, K), R = 1 << m, G = (1 << LP) >> m;
  const uint32_t rtl = (__brev((uint32_t)tl) >> (32 - (S - LP))) << (LP - m);
  uint32_t *base = xw + wx<S>((int)rtl);
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    const uint32_t rrr = (__brev((uint32_t)rr) >> (32 - m)) << (S - m);   // compile-time; wx is additive over rrr | rtl
    uint32_t *dst = base + wx<S>((int)rrr);
    auto at = [&](int e) { return (uint32_t)w[(int)(__brev((uint32_t)e) >> (32 - (LP - m))) * R + rr]; };   // bin e of the run <- group rev(e)
    if constexpr (G == 8) {
      *reinterpret_cast<uint4 *>(dst) = make_uint4(at(0), at(1), at(2), at(3));
      *reinterpret_cast<uint4 *>(dst + 4) = make_uint4(at(4), at(5), at(6), at(7));
    } else if constexpr (G == 4) *reinterpret_cast<uint4 *>(dst) = make_uint4(at(0), at(1), at(2), at(3));
    else if constexpr (G == 2) *reinterpret_cast<uint2 *>(dst) = make_uint2(at(0), at(1));
    else *dst = (uint32_t)w[rr];
  }
}

template <int S, int LP, bool TW3 = false>
__global__ void __launch_bounds__(FGeo<S, LP>::THREADS, fast_min_blocks<S, LP>()) fft_fixed_fast_kernel(const FastArgs a) {
  using Gm = FGeo<S, LP>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS;
  static_assert(wx<S>(N - 1) < Gm::ROW, "the padded output row fits the chirp's shared memory");
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ int loud[CPC];
  if (threadIdx.x < CPC) loud[threadIdx.x] = 0;
  __syncthreads();
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *xb = smem + (size_t)c * Gm::ROW, *xm = xb + Gm::XB;
  const uint32_t *xo = xb + wx<S>(tl);
  const uint32_t stride = gridDim.x * CPC;
  const GenRow gr = {0u, 0u};
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const uint32_t row = row0 + c < a.n_rows ? row0 + c : a.n_rows - 1;   // a chirp slot past the last row recomputes the last row
    const uint32_t *grow = a.in + (size_t)row * N;
    uint32_t pre[16];
    if (LP == 4) load_beats<S, LP>(a, grow, gr, tl, 0, pre);
    int w[PTS];
    fast_passes<S, LP, false, true, TW3>(a, c, tl, pre, grow, gr, xb, xm, loud, nullptr, w);   // ends with a barrier: the exchange buffer is free
    if (a.l2_prefetch && tl == 0 && row0 + stride + c < a.n_rows) {
      const uint32_t *nxt = a.in + (size_t)(row0 + stride + c) * N;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(nxt), "r"(4 * N) : "memory");
    }
    fast_scatter_words<S, LP>(tl, w, xb);
    chirp_sync<S>(c);
    uint32_t *gout = a.out + (size_t)row * N + tl;
#pragma unroll
    for (int j = 0; j < PTS; ++j) gout[j * TPC] = xo[wx<S>(j * TPC)];
    chirp_sync<S>(c);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// GOS / OS-CFAR (CFARCoreWithLis.scala:13-245 with cfarAlgorithm = 1): each side's reference cells are kept sorted by the
// reference's LinearSorter and the threshold is built from the indexLagg-th / indexLead-th largest cell (:136-150); the
// threshold is 0 unless both windows are full (:172-173).  Both sides are the same function of the window start,
//   K_k[s] = k-th largest of x[s .. s+W-1],   lagging value of bin n = K[n+G+1], leading value = K[n-G-W],
// so one array per distinct k serves the whole row.  It is computed by rank counting, one thread per group of 8 neighbouring
// cells: cell i is the k-th largest of window s iff exactly k-1 cells of the window beat it (ties go to the lower index).
// Sliding the window from s to s-1 adds x[s-1] and drops x[s+W-1], so the rank of every cell that stays moves by two
// comparisons; the 8 cells of a thread look at the same window in the same step and share its two loads.  About 6 W
// instructions per cell instead of 16 W for the bitwise search of the generic kernel, and no shared-memory traffic to speak of.
struct GosArgs { int cmode, linear, kl1, kr1; };   // kl1 = indexLagg - 1, kr1 = indexLead - 1

template <int S> struct GGeo {
  static constexpr int N = 1 << S;
  static constexpr int WM = 128;                                           // margin on both sides of the row (>= windowCells)
  static constexpr int XCELLS = (N + 2 * WM) + ((N + 2 * WM) >> 4) + 8;    // words of the padded cell array ...
  static constexpr int XBG = XCELLS > FGeo<S>::XB ? XCELLS : FGeo<S>::XB;  // ... which takes the place of the FFT exchange buffer
  static constexpr int KM = 128;                                           // int16 slots in front of each K array (window starts < 0)
  static constexpr int ROW = XBG + FGeo<S>::XM + N + KM;                   // + int16 magnitudes + two int16 K arrays with their margins
};

// offset of cell (16 tl + o) in the padded array, relative to the thread's base X32 + 17 tl (one pad word per 16 cells:
// lanes 16 cells apart are 17 words apart, conflict-free)
__host__ __device__ constexpr int gpad(int o) { return (o + 128) + ((o + 128) >> 4); }

// The compiler turns `r += a >= b` into add + conditional move and merges the conditional stores into select chains (9
// instructions per cell and step); written as predicated PTX it is compare + predicated add / store (6 per cell and step).
__device__ __forceinline__ void inc_if_ge(int &r, int a, int b) { asm("{.reg .pred p; setp.ge.s32 p, %1, %2; @p add.s32 %0, %0, 1;}" : "+r"(r) : "r"(a), "r"(b)); }
__device__ __forceinline__ void inc_if_gt(int &r, int a, int b) { asm("{.reg .pred p; setp.gt.s32 p, %1, %2; @p add.s32 %0, %0, 1;}" : "+r"(r) : "r"(a), "r"(b)); }
__device__ __forceinline__ void dec_if_gt(int &r, int a, int b) { asm("{.reg .pred p; setp.gt.s32 p, %1, %2; @p add.s32 %0, %0, -1;}" : "+r"(r) : "r"(a), "r"(b)); }
__device__ __forceinline__ void sts16_if_eq(int r, int k, uint32_t saddr, int v) {
  asm volatile("{.reg .pred p; setp.eq.s32 p, %0, %1; @p st.shared.u16 [%2], %3;}" :: "r"(r), "r"(k), "r"(saddr), "h"((short)v) : "memory");
}

template <bool TWO>
__device__ __forceinline__ void gos_step(int (&r)[8], const int (&xp)[8], int i0, int e_lo, int e_hi, int vL, int vR, int kl, int kr,
                                         uint32_t kl_addr, uint32_t kr_addr) {
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    if (e < e_lo || e > e_hi) continue;
    sts16_if_eq(r[e], kl, kl_addr, i0 + e);
    if (TWO) sts16_if_eq(r[e], kr, kr_addr, i0 + e);
    inc_if_ge(r[e], vL, xp[e]);
    dec_if_gt(r[e], vR, xp[e]);
  }
}

// K arrays of the 8 cells i0 .. i0+7 (i0 = 16 tl + o0) over all the windows that hold them
template <int S, bool TWO>
__device__ __forceinline__ void gos_group(const int *xt, int tl, int o0, int W, int kl1, int kr1, short *KL, short *KR) {
  constexpr int N = 1 << S;
  int xp[8], r[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { xp[e] = xt[gpad(o0 + e)]; r[e] = 0; }
  // cells after i_e inside its first window [i_e, i_e + W - 1] that beat it
#pragma unroll
  for (int d = 1; d < 8; ++d) {
    const int v = xt[gpad(o0 + d)];
#pragma unroll
    for (int e = 0; e < 8; ++e) if (e < d) inc_if_gt(r[e], v, xp[e]);
  }
  for (int d = 8; d < W; ++d) {
    const int v = xt[gpad(o0 + d)];
#pragma unroll
    for (int e = 0; e < 8; ++e) inc_if_gt(r[e], v, xp[e]);
  }
#pragma unroll
  for (int q = 0; q < 7; ++q) {
    const int v = xt[gpad(o0 + W + q)];
#pragma unroll
    for (int e = 0; e < 8; ++e) if (e > q) inc_if_gt(r[e], v, xp[e]);
  }
  // diagonal s: window start 16 tl + o0 - s; cell e is at step t = s + e of its own W windows
  const int a0 = 16 * tl + o0;
  const uint32_t kl_base = (uint32_t)__cvta_generic_to_shared(KL), kr_base = (uint32_t)__cvta_generic_to_shared(KR);
  auto diag = [&](int s, int e_lo, int e_hi) {
    const int oa = o0 - s, aidx = a0 - s;
    const int vL = xt[gpad(oa - 1)], vR = xt[gpad(oa + W - 1)];
    const bool valid = (unsigned)aidx <= (unsigned)(N - W);
    gos_step<TWO>(r, xp, a0, e_lo, e_hi, vL, vR, valid ? kl1 : -1, valid ? kr1 : -1, kl_base + 2u * (uint32_t)(valid ? aidx : 0),
                  kr_base + 2u * (uint32_t)(valid ? aidx : 0));
  };
#pragma unroll
  for (int s = -7; s < 0; ++s) diag(s, -s, 7);
  for (int s = 0; s <= W - 8; ++s) diag(s, 0, 7);
#pragma unroll
  for (int q = 1; q < 8; ++q) diag(W - 8 + q, 0, 7 - q);
}

// Packed form of the same rank counting, for rows whose cells all lie in 0 .. 32766 (every row that is not close to full scale;
// the others take gos_group above).  Two neighbouring cells share a register, the cells negated, and
//   [v > x]  = max(min(v - x, 1), 0)      [v >= x] = max(min(v + 1 - x, 1), 0)
// are ONE VIADDMNMX.S16x2.RELU for both cells (no 16-bit overflow: v + 1 <= 32767, -x >= -32766); its third operand masks the
// cells that have not started / are done (0x0001, 0x10000).  A half-register holds 8 * rank + e (e = 0 .. 7, the cell's place
// in the group), the update 8 * ([vL >= x] - [vR > x]) is two IMADs with a run-time +-8 (FMA pipe, as in the butterflies).
// "Some cell of mine has rank k" is then a minimum: min over the pairs of (half - 8 k) mod 2^16 is below 8 iff there is one,
// and it is its e -- four VIADDMNMX.U16x2 (add + min in one), one more min across the halves, one LOP3 with predicate output
// and ONE guarded 16-bit store of the cell's index per step (a predicated-off STS still costs its LSU slot on B200, measured
// in scripts/ubench/pipes.cu: eight of them per step were the bound of an earlier version).  Windows that start in front of
// the row land in the margin of the K arrays instead of being tested for.  Per step and 8 cells: 16 ALU-pipe + 11 FMA-pipe
// instructions + 2 LDS + 1 STS, against 42 + 16 + 8 for the scalar form.
__device__ __forceinline__ uint32_t gos_dup(int v, uint32_t add) { return (uint32_t)v * 0x10001u + add; }
__device__ __forceinline__ void gos_acc(uint32_t &r, uint32_t d, int mul) { asm("mad.lo.s32 %0, %1, %2, %0;" : "+r"(r) : "r"(d), "r"(mul)); }
// if ((key & 0xfff8) == 0) K[slot] = i0 + key   (low half of key: 8 * (rank - k) + e of the best cell)
__device__ __forceinline__ void gos_store_if_match(uint32_t key, uint32_t saddr, int i0, int one) {
  uint32_t idx = (uint32_t)i0;
  gos_acc(idx, key, one);
  asm volatile("{.reg .pred p, q; .reg .b32 d; setp.ne.u32 q, %0, %0; lop3.or.b32 d|p, %0, 0xfff8, 0, 0xc0, q; @!p st.shared.u16 [%1], %2;}"
               :: "r"(key), "r"(saddr), "h"((short)idx) : "memory");
}

// 8 consecutive steps.  MODE 0: all 8 cells take part; 1: start-up, step u has the cells e >= 7 - u (u = 0 .. 6); 2: run-out,
// step u has the cells e <= 7 - u.  pl / pr point at the word of step 0's entering / leaving cell, later steps go down by one
// (the 8 cells of a chunk sit in one aligned group of 8, so there is no pad word between them); ka: K slot of step 0.
template <bool TWO, int MODE>
__device__ __forceinline__ void gos_chunk(uint32_t (&r)[4], const uint32_t (&nx)[4], const int *pl, const int *pr, uint32_t nkl, uint32_t nkr,
                                          uint32_t kal, uint32_t kar, int i0, int one, int p8, int m8) {
#pragma unroll
  for (int u = 0; u < (MODE == 1 ? 7 : 8); ++u) {
    const uint32_t vl1 = gos_dup(pl[-u], 0x10001u), vr = gos_dup(pr[-u], 0u);
    uint32_t ml = 0xFFFFFFFFu, mr = 0xFFFFFFFFu;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const bool lo = MODE == 0 || (MODE == 1 ? 2 * p >= 7 - u : 2 * p <= 7 - u);
      const bool hi = MODE == 0 || (MODE == 1 ? 2 * p + 1 >= 7 - u : 2 * p + 1 <= 7 - u);
      if (!lo && !hi) continue;
      const uint32_t live = r[p] | (lo ? 0u : 0xFFFFu) | (hi ? 0u : 0xFFFF0000u);   // a cell outside its windows never matches
      ml = __viaddmin_u16x2(live, (lo ? nkl & 0xFFFFu : 0u) | (hi ? nkl & 0xFFFF0000u : 0u), ml);
      if (TWO) mr = __viaddmin_u16x2(live, (lo ? nkr & 0xFFFFu : 0u) | (hi ? nkr & 0xFFFF0000u : 0u), mr);
      const uint32_t ones = (lo ? 1u : 0u) | (hi ? 0x10000u : 0u);
      gos_acc(r[p], __viaddmin_s16x2_relu(vl1, nx[p], ones), p8);
      gos_acc(r[p], __viaddmin_s16x2_relu(vr, nx[p], ones), m8);
    }
    gos_store_if_match(__vminu2(ml, ml >> 16), kal - 2u * u, i0, one);
    if (TWO) gos_store_if_match(__vminu2(mr, mr >> 16), kar - 2u * u, i0, one);
  }
}

template <int S, bool TWO, int O0>
__device__ __forceinline__ void gos_group_packed(const int *xt, int tl, int W, int kl1, int kr1, short *KL, short *KR, int one, int m1) {
  constexpr int B0 = gpad(O0);
  const int p8 = one * 8, m8 = m1 * 8;
  uint32_t nx[4], r[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    nx[p] = __byte_perm((uint32_t)(-xt[B0 + 2 * p]), (uint32_t)(-xt[B0 + 2 * p + 1]), 0x5410);
    r[p] = (uint32_t)(2 * p) | ((uint32_t)(2 * p + 1) << 16);
  }
  // cells after i_e inside its first window [i_e, i_e + W - 1] that beat it
#pragma unroll
  for (int d = 1; d < 8; ++d) {
    const uint32_t v = gos_dup(xt[B0 + d], 0u);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const uint32_t ones = (2 * p < d ? 1u : 0u) | (2 * p + 1 < d ? 0x10000u : 0u);
      if (ones) gos_acc(r[p], __viaddmin_s16x2_relu(v, nx[p], ones), p8);
    }
  }
  for (int j = 1; j < (W >> 3); ++j) {
    const int idx = O0 + 128 + 8 * j;
    const int *pv = xt + idx + (idx >> 4);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint32_t v = gos_dup(pv[u], 0u);
#pragma unroll
      for (int p = 0; p < 4; ++p) gos_acc(r[p], __viaddmin_s16x2_relu(v, nx[p], 0x10001u), p8);
    }
  }
  {
    const int *pv = xt + gpad(O0 + W);
#pragma unroll
    for (int q = 0; q < 7; ++q) {
      const uint32_t v = gos_dup(pv[q], 0u);
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const uint32_t ones = (2 * p > q ? 1u : 0u) | (2 * p + 1 > q ? 0x10000u : 0u);
        if (ones) gos_acc(r[p], __viaddmin_s16x2_relu(v, nx[p], ones), p8);
      }
    }
  }
  // step s looks at the window that starts at a0 - s: s = -7 .. -1 start-up, 0 .. W - 9 all cells, W - 8 .. W - 1 run-out
  const int a0 = 16 * tl + O0;
  const uint32_t nkl = ((uint32_t)(-8 * kl1) & 0xFFFFu) * 0x10001u, nkr = ((uint32_t)(-8 * kr1) & 0xFFFFu) * 0x10001u;   // -8 k in both halves
  const uint32_t kl0 = (uint32_t)__cvta_generic_to_shared(KL) + 2u * (uint32_t)a0, kr0 = (uint32_t)__cvta_generic_to_shared(KR) + 2u * (uint32_t)a0;
  gos_chunk<TWO, 1>(r, nx, xt + B0 + 6, xt + gpad(O0 + W) + 6, nkl, nkr, kl0 + 14u, kr0 + 14u, a0, one, p8, m8);
  for (int j = 0; j < (W >> 3) - 1; ++j) {
    const int il = O0 + 128 - 8 - 8 * j, ir = il + W;
    gos_chunk<TWO, 0>(r, nx, xt + il + (il >> 4) + 7, xt + ir + (ir >> 4) + 7, nkl, nkr, kl0 - 16u * (uint32_t)j, kr0 - 16u * (uint32_t)j, a0, one, p8, m8);
  }
  {
    const int il = O0 + 128 - W;
    const uint32_t ko = 2u * (uint32_t)(W - 8);
    gos_chunk<TWO, 2>(r, nx, xt + il + (il >> 4) + 7, xt + B0 + 7, nkl, nkr, kl0 - ko, kr0 - ko, a0, one, p8, m8);
  }
}

template <int S>
__global__ void __launch_bounds__(FGeo<S>::THREADS, 1024 / FGeo<S>::THREADS) chain_fixed_fast_gos_kernel(const FastArgs a, const GosArgs ga) {
  using Gm = FGeo<S>;
  using Gg = GGeo<S>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS;
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ int loud[CPC];
  if (threadIdx.x < CPC) loud[threadIdx.x] = 0;
  __syncthreads();
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *xb = smem + (size_t)c * Gg::ROW, *xm = xb + Gg::XBG;
  const short *X = reinterpret_cast<const short *>(xm);
  short *KL = reinterpret_cast<short *>(xm + Gm::XM) + Gg::KM, *KR = ga.kl1 == ga.kr1 ? KL : KL + N + Gg::KM;
  int *xt = reinterpret_cast<int *>(xb) + 17 * tl;
  const uint32_t stride = gridDim.x * CPC;
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const bool active = row0 + c < a.n_rows;
    const uint32_t row = active ? row0 + c : a.n_rows - 1;
    const uint32_t *grow = a.in + (size_t)row * N;
    uint32_t pre[PTS];
    const GenRow no_gen = {0u, 0u};   // (the generator is fused into the CA kernel only; a.gen_full is null here)
    load_beats<S, 4>(a, grow, no_gen, tl, 0, pre);
    fast_passes<S, 4>(a, c, tl, pre, grow, no_gen, xb, xm, loud);
    uint32_t wide = 0;
    {   // the thread's 16 cells as 32-bit words in the padded array (the exchange buffer is dead)
      const uint32_t *xq = xm + mhx<S>(16 * tl) / 2;   // the thread's 16 consecutive bins
      const uint4 v0 = *reinterpret_cast<const uint4 *>(xq), v1 = *reinterpret_cast<const uint4 *>(xq + 4);
      const uint32_t wd[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) { xt[gpad(2 * k)] = sext16((int)wd[k]); xt[gpad(2 * k + 1)] = hi16s(wd[k]); }
#pragma unroll
      for (int k = 0; k < 8; ++k) wide |= (wd[k] + 0x00010001u) | wd[k];   // bit 15 of a half: the cell is negative or 32767
    }
    if (__syncthreads_or((int)(wide & 0x80008000u))) {   // some cell of the CTA's rows is outside 0 .. 32766: scalar ranks
      for (int h = 0; h < 2; ++h) {
        if (ga.kl1 == ga.kr1) gos_group<S, false>(xt, tl, 8 * h, a.W, ga.kl1, ga.kr1, KL, KR);
        else gos_group<S, true>(xt, tl, 8 * h, a.W, ga.kl1, ga.kr1, KL, KR);
      }
    } else if (ga.kl1 == ga.kr1) {
      gos_group_packed<S, false, 0>(xt, tl, a.W, ga.kl1, ga.kr1, KL, KR, a.one, a.m1);
      gos_group_packed<S, false, 8>(xt, tl, a.W, ga.kl1, ga.kr1, KL, KR, a.one, a.m1);
    } else {
      gos_group_packed<S, true, 0>(xt, tl, a.W, ga.kl1, ga.kr1, KL, KR, a.one, a.m1);
      gos_group_packed<S, true, 8>(xt, tl, a.W, ga.kl1, ga.kr1, KL, KR, a.one, a.m1);
    }
    chirp_sync<S>(c);
    uint32_t *gout = a.out + (size_t)row * N;
#pragma unroll 4
    for (int j = 0; j < PTS; ++j) {
      const int n = tl + j * TPC, W = a.W, G = a.G;
      const bool both = (n + G + W <= N - 1) && (n - G - W >= 0);
      const int L = X[mhx<S>(KL[min(n + G + 1, N - W)])], R = X[mhx<S>(KR[max(n - G - W, 0)])];   // the K arrays hold the cell's index
      int t = ga.cmode == 0 ? ((R + L) >> 1) : (ga.cmode == 1 ? max(L, R) : min(L, R));
      t = both ? t : 0;                                                  // CFARCoreWithLis.scala:172-173
      const int thr = ga.linear ? t * a.scaler : t + a.scaler;
      const int cut = X[mhx<S>(n)];
      const int pk = cut > thr ? 1 : 0;
      gout[n] = ((uint32_t)thr & a.thr_mask) * (uint32_t)a.out_mul + (uint32_t)(n << 1) + (uint32_t)pk;
      if (pk && a.det_list != nullptr && active) {
        const uint32_t slot = atomicAdd(a.det_count, 1u);
        if (slot < a.det_cap) {
          Detection d; d.chirp = a.row_base + row; d.bin = (uint32_t)n; d.threshold = (int)((uint32_t)thr << a.thr_sext) >> a.thr_sext; d.cut = cut;
          a.det_list[slot] = d;
        }
      }
    }
    chirp_sync<S>(c);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Combined-window ordered statistic (CORE_OS: the golden cfarOS, CFARUtils.scala:159-196; BASELINE configs[3] reading (B)) by radix
// selection on bit planes.  The row's biased magnitudes (x + 2^15, 16 bits) are transposed into 16 bit planes of N bits each (one warp
// ballot per plane and 32 bins); the reference cells of a bin -- R <= 32 leading and R lagging ones -- are then two 32-bit masks cut out
// of a plane by funnel shifts, and the k-th largest of the 2 R cells is found from the top plane down: with the candidates' bit in hand,
// c = popc(cand & plane) says whether the statistic has the bit set (c >= k: keep those) or not (k -= c, keep the others).  About 12
// instructions per plane and bin, independent of R, against 2 R compares per plane for the bitwise rank search of the run-time-parameter
// kernel.  A thread owns 16 neighbouring bins, 8 at a time: their windows are 8 consecutive bit offsets of the same three plane words.
// The planes are built by a 16 x 16 bit transpose in registers per thread; planes in which all cells of the chirp agree are skipped.
#ifndef RSP_OS_SPARSE
#define RSP_OS_SPARSE 1   // A/B knob: 0 = every window visits all the planes that are not uniform over the chirp
#endif
constexpr bool SPARSE = RSP_OS_SPARSE != 0;
template <int S> struct OGeo {
  static constexpr int N = 1 << S;
  static constexpr int MW = 3, MB = 6;               // margin words in front of / behind each plane (windows that start before or end after the row)
  static constexpr int PW = N / 32 + MW + MB;        // words per plane (odd: the planes of one word index fall into distinct banks)
  static_assert(PW & 1, "odd plane stride");
  static_assert(16 * PW <= FGeo<S>::XB, "the bit planes fit the exchange buffer");
  static constexpr int KW = N;                       // extra words of the (A) kernel: two int16 K arrays
  static constexpr int NZ = 16 * PW;                 // word offset (from the start of the exchange buffer) of the per-block plane masks
  static constexpr int NZM = 16;                     // zero entries behind them (windows that end after the row)
  static_assert(NZ + (N / 16 + NZM + 1) / 2 <= FGeo<S>::XB, "the plane masks fit the exchange buffer behind the planes");
};

// Peak grouping on the core that reads its neighbours from the guard registers of both sides (CFARCoreWithLis.scala:191-192): a peak must also exceed x[n + 1] (read only when guardCells equals the generator's
// guardWindowSize, else 0) and x[n - 1]; with guardCells = 0 both neighbours read 0.  X: the chirp's magnitudes (mhx layout).
template <int S>
__device__ __forceinline__ int lis_peak_grouping(const FastArgs &a, const short *X, int n, int cut, int pk) {
  constexpr int N = 1 << S;
  int left = 0, right = 0;
  if (a.G != 0) {
    if (a.pg_full > 0 && n + 1 < N) left = X[mhx<S>(n + 1)];
    if (n >= 1) right = X[mhx<S>(n - 1)];
  }
  return (pk != 0 && cut > left && cut > right) ? 1 : 0;
}

// 16 x 16 bit-matrix transpose in 8 registers (word k = rows 2k | 2k+1 << 16): block swaps of size 8, 4, 2, 1.  In: row r = the 16-bit
// value of bin r; out: row c = bit c of the 16 bins (bin r at bit r).  ~100 instructions per 16 bins, against 4 per bin and plane for
// warp ballots (tests/test_kernel_layout.py holds the same algorithm in Python).
__device__ __forceinline__ void bit_transpose16(uint32_t (&A)[8]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) { const uint32_t t = ((A[k] >> 8) ^ A[k + 4]) & 0x00FF00FFu; A[k + 4] ^= t; A[k] ^= t << 8; }
#pragma unroll
  for (int q = 0; q < 4; ++q) { const int k = (q & 1) | ((q & 2) << 1); const uint32_t t = ((A[k] >> 4) ^ A[k + 2]) & 0x0F0F0F0Fu; A[k + 2] ^= t; A[k] ^= t << 4; }
#pragma unroll
  for (int k = 0; k < 8; k += 2) { const uint32_t t = ((A[k] >> 2) ^ A[k + 1]) & 0x33333333u; A[k + 1] ^= t; A[k] ^= t << 2; }
#pragma unroll
  for (int k = 0; k < 8; ++k) { const uint32_t t = ((A[k] >> 1) ^ (A[k] >> 16)) & 0x5555u; A[k] ^= t * 0x10002u; }
}

// The thread's 16 neighbouring magnitudes (biased to offset binary: order preserved) -> its halfword of each of the 16 planes.
// uni[0] / uni[1] collect the OR / AND of all cells of the chirp: a plane in which every cell has the same bit cannot change any
// candidate set (all candidates have the bit: c >= k as long as k <= candidates; none has it: c = 0) and is skipped by the selection.
// nz[t] = the planes in which at least one of thread t's 16 cells has its bit set: a selection step over a window whose cells all have
// a 0 in the plane leaves every candidate set as it is (c = 0 < k), so the selection ORs the masks of the blocks its windows touch and
// visits only those planes -- on radar rows (a few targets over a noise floor of a few LSB) most windows see three or four planes.
template <int S>
__device__ __forceinline__ void os_build_planes(const uint32_t *xm, int tl, uint32_t *bp, uint32_t *uni, uint16_t *nz) {
  constexpr int PW = OGeo<S>::PW;
  const uint32_t *xq = xm + mhx<S>(16 * tl) / 2;
  const uint4 v0 = *reinterpret_cast<const uint4 *>(xq), v1 = *reinterpret_cast<const uint4 *>(xq + 4);
  uint32_t A[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
  uint32_t orv = 0, andv = 0xFFFFFFFFu;
#pragma unroll
  for (int k = 0; k < 8; ++k) { A[k] ^= 0x80008000u; orv |= A[k]; andv &= A[k]; }
  nz[tl] = (uint16_t)(orv | (orv >> 16));
  if (tl < OGeo<S>::NZM) nz[FGeo<S>::TPC + tl] = 0;
  orv = __reduce_or_sync(0xffffffffu, orv); andv = __reduce_and_sync(0xffffffffu, andv);
  if ((tl & 31) == 0) { atomicOr(&uni[0], (orv | (orv >> 16)) & 0xFFFFu); atomicAnd(&uni[1], andv & (andv >> 16) & 0xFFFFu); }
  bit_transpose16(A);
  uint16_t *bph = reinterpret_cast<uint16_t *>(bp) + tl;
#pragma unroll
  for (int b = 0; b < 16; ++b) bph[b * 2 * PW] = (uint16_t)(A[b >> 1] >> (16 * (b & 1)));
}

// planes that can matter for windows over the bins lo .. hi (clamped to the row in front; the zero entries take care of the end)
__device__ __forceinline__ uint32_t os_local_planes(const uint16_t *nz, uint32_t act, int lo, int hi) {
  uint32_t m = 0;
  for (int b = max(lo, 0) >> 4; b <= (hi >> 4); ++b) m |= nz[b];
  return act & m;
}

// 8 consecutive window pairs of the combined-window statistic: leading words start at bit s0 + j, lagging words at bit s1 + j (j < 8) of
// every plane; m0 / m1 mask the cells that belong to the window; kk0 = rank (1-based, descending); act = planes that are not uniform.
// Leaves the candidate masks of the statistic in cl / cr.
template <int S>
__device__ __forceinline__ void os_select8(const uint32_t *bp, uint32_t act, int s0, int s1, uint32_t m0, uint32_t m1, int kk0, uint32_t (&cl)[8], uint32_t (&cr)[8]) {
  constexpr int PW = OGeo<S>::PW;
  const uint32_t *p0 = bp + (s0 >> 5), *p1 = bp + (s1 >> 5);   // arithmetic shift: the margin words take the negative starts
  const int sh0 = s0 & 31, sh1 = s1 & 31;
  int kk[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { cl[j] = m0; cr[j] = m1; kk[j] = kk0; }
  while (act) {
    const int b = 31 - __clz((int)act);
    act ^= 1u << b;
    const uint32_t a0 = p0[b * PW], a1 = p0[b * PW + 1], a2 = p0[b * PW + 2];
    const uint32_t c0 = p1[b * PW], c1 = p1[b * PW + 1], c2 = p1[b * PW + 2];
    const uint32_t l0 = __funnelshift_r(a0, a1, sh0), l1 = __funnelshift_r(a1, a2, sh0);   // bits s0 .. s0 + 63
    const uint32_t r0 = __funnelshift_r(c0, c1, sh1), r1 = __funnelshift_r(c1, c2, sh1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t wl = __funnelshift_r(l0, l1, j), wr = __funnelshift_r(r0, r1, j);
      const uint32_t ml = cl[j] & wl, mr = cr[j] & wr;
      const int t = kk[j] - __popc(ml) - __popc(mr);
      const bool take = t <= 0;                      // at least kk candidates have the bit: the statistic has it too
      const uint32_t tm = take ? 0u : 0xFFFFFFFFu;
      cl[j] = ml ^ (tm & cl[j]);                     // take: the candidates with the bit; else: those without
      cr[j] = mr ^ (tm & cr[j]);
      kk[j] = take ? kk[j] : t;
    }
  }
}

// The same for ONE window of W <= 32 NWD cells per start (the hardware GOS core, CFARCoreWithLis.scala:136-143): 8 consecutive starts
// s0 + j; mlast masks the cells of the last word.  Leaves the candidate masks in cw[j][0 .. NWD).
template <int S, int NWD>
__device__ __forceinline__ void gos_select8(const uint32_t *bp, uint32_t act, int s0, uint32_t mlast, int kk0, uint32_t (&cw)[8][NWD]) {
  constexpr int PW = OGeo<S>::PW;
  const uint32_t *p0 = bp + (s0 >> 5);
  const int sh0 = s0 & 31;
  int kk[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    kk[j] = kk0;
#pragma unroll
    for (int w = 0; w < NWD; ++w) cw[j][w] = w == NWD - 1 ? mlast : 0xFFFFFFFFu;
  }
  while (act) {
    const int b = 31 - __clz((int)act);
    act ^= 1u << b;
    uint32_t raw[NWD + 2], al[NWD + 1];
#pragma unroll
    for (int w = 0; w < NWD + 2; ++w) raw[w] = p0[b * PW + w];
#pragma unroll
    for (int w = 0; w < NWD + 1; ++w) al[w] = __funnelshift_r(raw[w], raw[w + 1], sh0);   // bits s0 .. s0 + 32 (NWD + 1) - 1
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint32_t m[NWD];
      int t = kk[j];
#pragma unroll
      for (int w = 0; w < NWD; ++w) { m[w] = cw[j][w] & __funnelshift_r(al[w], al[w + 1], j); t -= __popc(m[w]); }
      const bool take = t <= 0;
      const uint32_t tm = take ? 0u : 0xFFFFFFFFu;
#pragma unroll
      for (int w = 0; w < NWD; ++w) cw[j][w] = m[w] ^ (tm & cw[j][w]);
      kk[j] = take ? kk[j] : t;
    }
  }
}

template <int S, bool TW3 = false>
__global__ void __launch_bounds__(FGeo<S>::THREADS, 1024 / FGeo<S>::THREADS) chain_fixed_fast_os_kernel(const FastArgs a, const GosArgs ga) {
  using Gm = FGeo<S>;
  using Go = OGeo<S>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS;
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ int loud[CPC];
  __shared__ uint32_t uni[CPC][2];
  if (threadIdx.x < CPC) loud[threadIdx.x] = 0;
  __syncthreads();
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *xb = smem + (size_t)c * Gm::ROW, *xm = xb + Gm::XB;
  const short *X = reinterpret_cast<const short *>(xm);
  uint32_t *bp = xb + Go::MW;                         // plane b, word w (bins 32 w .. 32 w + 31): bp[b * PW + w]
  uint16_t *nz = reinterpret_cast<uint16_t *>(xb + Go::NZ);
  const int R = a.W, G = a.G;
  const uint32_t mR = R >= 32 ? 0xFFFFFFFFu : ((1u << R) - 1u);
  const int kk0 = 2 * R - (3 * R) / 2;                // ascending index floor(1.5 R) of 2 R cells = the (2 R - floor(1.5 R))-th largest
  const uint32_t stride = gridDim.x * CPC;
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const bool active = row0 + c < a.n_rows;
    const uint32_t row = active ? row0 + c : a.n_rows - 1;
    const uint32_t *grow = a.in + (size_t)row * N;
    uint32_t pre[PTS];
    const GenRow no_gen = {0u, 0u};
    if (tl == 0) { uni[c][0] = 0u; uni[c][1] = 0xFFFFu; }
    if (a.cfar_only) {   // stored profiles: the cells straight into the magnitude array (bin n = tl + j TPC: warp-contiguous loads)
      short *Xw = reinterpret_cast<short *>(xm);
#pragma unroll
      for (int j = 0; j < PTS; ++j) Xw[mhx<S>(tl + j * TPC)] = (short)__ldg(&grow[tl + j * TPC]);
      chirp_sync<S>(c);
    } else {
      load_beats<S, 4>(a, grow, no_gen, tl, 0, pre);
      fast_passes<S, 4, false, false, TW3>(a, c, tl, pre, grow, no_gen, xb, xm, loud);   // ends with a barrier: magnitudes in X, the exchange buffer is free
    }
    os_build_planes<S>(xm, tl, bp, uni[c], nz);
    chirp_sync<S>(c);
    const uint32_t act = uni[c][0] ^ uni[c][1];       // planes in which the chirp's cells differ
    uint32_t *gout = a.out + (size_t)row * N + PTS * tl;
#pragma unroll 1
    for (int h = 0; h < PTS / 8; ++h) {
      const int n0 = PTS * tl + 8 * h;
      uint32_t cl[8], cr[8];
      os_select8<S>(bp, SPARSE ? os_local_planes(nz, act, n0 - G - R, n0 + G + R + 7) : act, n0 - G - R, n0 + G + 1, mR, mR, kk0, cl, cr);
      uint32_t ow[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = n0 + j;
        const bool both = (n + G + R <= N - 1) && (n - G - R >= 0);
        // every remaining candidate equals the statistic: read one back from the magnitudes
        int pos = cl[j] != 0 ? (n - G - R) + __ffs((int)cl[j]) - 1 : (n + G + 1) + __ffs((int)cr[j]) - 1;
        pos = min(max(pos, 0), N - 1);
        const int t = X[mhx<S>(pos)];
        int thr = ga.linear ? sar(t * a.scaler, a.thr_sh) : t * a.t_mul + a.sc_pre;   // (binary points at run time, see the GOS kernel)
        thr = both ? thr : 0;                         // the golden pads the edges with 0 (linear and log)
        const int cut = X[mhx<S>(n)];
        const int pk = (both && (cut * a.ncut_mul + thr) < 0) ? 1 : 0;   // (no peak grouping in the cfarOS golden: refused by the host)
        thr = sar(thr, a.out_sh);
        ow[j] = ((uint32_t)thr & a.thr_mask) * (uint32_t)a.out_mul + (uint32_t)(n << 1) + (uint32_t)pk;
        if (pk && a.det_list != nullptr && active) {
          const uint32_t slot = atomicAdd(a.det_count, 1u);
          if (slot < a.det_cap) {
            Detection d; d.chirp = a.row_base + row; d.bin = (uint32_t)n; d.threshold = (int)((uint32_t)thr << a.thr_sext) >> a.thr_sext; d.cut = cut;
            a.det_list[slot] = d;
          }
        }
      }
      if (a.out != nullptr) {
        *reinterpret_cast<uint4 *>(gout + 8 * h) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
        *reinterpret_cast<uint4 *>(gout + 8 * h + 4) = make_uint4(ow[4], ow[5], ow[6], ow[7]);
      }
    }
    chirp_sync<S>(c);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// The hardware GOS core (CFARCoreWithLis, reading (A) of BASELINE configs[3]) on the same bit planes: K_k[s] = k-th largest of the W
// cells starting at s is one radix selection per window start over NWD = ceil(W / 32) mask words (W <= 128), independent of k and -- per
// mask word -- of W; the thresholds then read K[n + G + 1] and K[n - G - W] as in chain_fixed_fast_gos_kernel.  TWO: indexLagg and
// indexLead differ, two selections.  No restriction on the cells' range (the bias makes them unsigned) or on W being a multiple of 8.
// PG: peak grouping (the registers' other setting) as an instantiation of its own, so the default one is what it was.
template <int S, int NWD, bool TW3, bool PG = false>
__global__ void __launch_bounds__(FGeo<S>::THREADS, (NWD >= 3 ? 512 : 1024) / FGeo<S>::THREADS >= 1 ? (NWD >= 3 ? 512 : 1024) / FGeo<S>::THREADS : 1)
chain_fixed_fast_gosb_kernel(const FastArgs a, const GosArgs ga) {
  using Gm = FGeo<S>;
  using Go = OGeo<S>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS, CROW = Gm::ROW + Go::KW;
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ int loud[CPC];
  __shared__ uint32_t uni[CPC][2];
  if (threadIdx.x < CPC) loud[threadIdx.x] = 0;
  __syncthreads();
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *xb = smem + (size_t)c * CROW, *xm = xb + Gm::XB;
  const short *X = reinterpret_cast<const short *>(xm);
  const bool two = ga.kl1 != ga.kr1;   // indexLagg and indexLead differ: two selections, two K arrays
  short *KL = reinterpret_cast<short *>(xm + Gm::XM), *KR = two ? KL + N : KL;
  uint32_t *bp = xb + Go::MW;
  uint16_t *nz = reinterpret_cast<uint16_t *>(xb + Go::NZ);
  const int W = a.W, G = a.G;
  const uint32_t mlast = (W & 31) ? ((1u << (W & 31)) - 1u) : 0xFFFFFFFFu;
  const uint32_t stride = gridDim.x * CPC;
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const bool active = row0 + c < a.n_rows;
    const uint32_t row = active ? row0 + c : a.n_rows - 1;
    const uint32_t *grow = a.in + (size_t)row * N;
    uint32_t pre[PTS];
    const GenRow no_gen = {0u, 0u};
    if (tl == 0) { uni[c][0] = 0u; uni[c][1] = 0xFFFFu; }
    if (a.cfar_only) {   // stored profiles (rsp_cfar_run): see chain_fixed_fast_os_kernel
      short *Xw = reinterpret_cast<short *>(xm);
#pragma unroll
      for (int j = 0; j < PTS; ++j) Xw[mhx<S>(tl + j * TPC)] = (short)__ldg(&grow[tl + j * TPC]);
      chirp_sync<S>(c);
    } else {
      load_beats<S, 4>(a, grow, no_gen, tl, 0, pre);
      fast_passes<S, 4, false, false, TW3>(a, c, tl, pre, grow, no_gen, xb, xm, loud);
    }
    os_build_planes<S>(xm, tl, bp, uni[c], nz);
    chirp_sync<S>(c);
    const uint32_t act = uni[c][0] ^ uni[c][1];
#pragma unroll 1
    for (int h = 0; h < (two ? 4 : 2); ++h) {          // window starts 16 tl + 8 (h & 1) + j; h >= 2: the leading statistic
      const int s0 = PTS * tl + 8 * (h & 1);
      uint32_t cw[8][NWD];
      gos_select8<S, NWD>(bp, SPARSE ? os_local_planes(nz, act, s0, s0 + W + 6) : act, s0, mlast, (h < 2 ? ga.kl1 : ga.kr1) + 1, cw);
      short *K = h < 2 ? KL : KR;
      uint32_t kv[4];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        int pos = s0 + j;                               // every remaining candidate equals the statistic: read the first one back
        bool found = false;
#pragma unroll
        for (int w = 0; w < NWD; ++w) { if (!found && cw[j][w] != 0) { pos += 32 * w + __ffs((int)cw[j][w]) - 1; found = true; } }
        const uint32_t v = (uint16_t)X[mhx<S>(min(pos, N - 1))];
        if (j & 1) kv[j >> 1] |= v << 16; else kv[j >> 1] = v;
      }
      *reinterpret_cast<uint4 *>(&K[s0]) = make_uint4(kv[0], kv[1], kv[2], kv[3]);
    }
    chirp_sync<S>(c);
    uint32_t *gout = a.out + (size_t)row * N;
#pragma unroll 4
    for (int j = 0; j < PTS; ++j) {
      const int n = tl + j * TPC;
      const bool both = (n + G + W <= N - 1) && (n - G - W >= 0);
      const int L = KL[min(n + G + 1, N - W)], R = KR[max(n - G - W, 0)];
      int t = ga.cmode == 0 ? ((R + L) >> 1) : (ga.cmode == 1 ? max(L, R) : min(L, R));
      t = both ? t : 0;                                                  // CFARCoreWithLis.scala:172-173
      // (the binary-point form of fast_cfar_bin at run time: without binary points the shifts are 0, t_mul 1, ncut_mul -1, sc_pre = scaler)
      const int thr = ga.linear ? sar(t * a.scaler, a.thr_sh) : t * a.t_mul + a.sc_pre;
      const int cut = X[mhx<S>(n)];
      int pk = (cut * a.ncut_mul + thr) < 0 ? 1 : 0;
      if (PG) pk = lis_peak_grouping<S>(a, X, n, cut, pk);
      const int tf = sar(thr, a.out_sh);
      if (a.out != nullptr) gout[n] = ((uint32_t)tf & a.thr_mask) * (uint32_t)a.out_mul + (uint32_t)(n << 1) + (uint32_t)pk;
      if (pk && a.det_list != nullptr && active) {
        const uint32_t slot = atomicAdd(a.det_count, 1u);
        if (slot < a.det_cap) {
          Detection d; d.chirp = a.row_base + row; d.bin = (uint32_t)n; d.threshold = (int)((uint32_t)tf << a.thr_sext) >> a.thr_sext; d.cut = cut;
          a.det_list[slot] = d;
        }
      }
    }
    chirp_sync<S>(c);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// host side.  This file is compiled once per transform size (-DRSP_FAST_S=10 .. 14, see Makefile) so that the sizes build in
// parallel; the unit of the smallest size also holds the dispatcher.
#ifndef RSP_FAST_S
#error "compile with -DRSP_FAST_S=10..14"
#endif
static int fast_edge_iters(int S, int W, int G) { const int tpc = (1 << S) >> (S == 14 ? 5 : 4); return (W + G + tpc - 1) / tpc; }   // threads per chirp of the narrower geometry

// cp.async staging of the next row (FastArgs::stage).  Measured on B200 (scripts/gpu_stage_ab.sh, profiles/r01_l_stage_ab.txt): +1.9 %
// at N = 16384 with 16 points per thread (one CTA per SM, nobody else to hide the row's first loads), nothing or slightly negative
// below, so it is on from 2^14 points (and only with 16 points per thread: two 105 KB rows leave no room for a staging area);
// RSP_FAST_STAGE=0 turns it off, RSP_FAST_STAGE_MIN_S=s moves the limit (tuning knobs, read once)
static bool fast_stage_enabled(int S) {
  static int on = -1, min_s = 0;
  if (on < 0) {
    const char *e = std::getenv("RSP_FAST_STAGE"), *m = std::getenv("RSP_FAST_STAGE_MIN_S");
    on = (e != nullptr && e[0] == '0') ? 0 : 1;
    min_s = m != nullptr ? std::atoi(m) : 14;
  }
  return on != 0 && S >= min_s;
}

// points per thread at N = 16384: 32 (two CTAs per SM) unless RSP_FAST_LP14=4 asks for the one-CTA form (A/B knob, read once)
static int fast_lp_runtime(int S) {
  static int lp14 = -1;
  if (lp14 < 0) { const char *e = std::getenv("RSP_FAST_LP14"); lp14 = (e != nullptr && e[0] == '4') ? 4 : RSP_FAST_LP14; }
  return S == 14 ? lp14 : 4;
}

template <int S, int LP, int CMODE, bool LINEAR, int EJ, bool PG = false, bool TW3 = false, bool BPC = false>
static rsp_status launch_fast_inst_lp(const LaunchCtx &c, const FastArgs &fa_in, cudaStream_t st, const char **err) {
  using Gm = FGeo<S, LP>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  FastArgs fa = fa_in;
  if (LP != 4) fa.stage = 0;
  const bool stage = LP == 4 && fast_stage_enabled(S);
  const size_t smem = ((size_t)Gm::CPC * (Gm::ROW + (Gm::TWS == 0 && stage ? Gm::N : 0)) + Gm::TWS) * sizeof(uint32_t);
  auto kernel = chain_fixed_fast_kernel<S, LP, CMODE, LINEAR, EJ, PG, TW3, BPC>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_fixed_fast_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}
template <int S, int CMODE, bool LINEAR, int EJ>
static rsp_status launch_fast_inst(const LaunchCtx &c, const FastArgs &fa, bool pg, cudaStream_t st, const char **err) {
  constexpr int LPV = S == 14 ? 5 : 4;
  if (fa.tw3 != nullptr) {   // binPoint >= 2; its CFAR phase is the binary-point one.  Peak grouping is an instantiation of its own (as a run-time flag
                             // it cost the applications' configuration 6.5 %)
    if (pg) return launch_fast_inst_lp<S, LPV, CMODE, LINEAR, EJ, true, true, true>(c, fa, st, err);
    return launch_fast_inst_lp<S, LPV, CMODE, LINEAR, EJ, false, true, true>(c, fa, st, err);
  }
  if (fa.bpc) {   // binPoint 0 in the transform, binary points in the CFAR (RspChainVanillaSpec: input 0, threshold 3, scaler 6)
    if (pg) return launch_fast_inst_lp<S, LPV, CMODE, LINEAR, EJ, true, false, true>(c, fa, st, err);
    return launch_fast_inst_lp<S, LPV, CMODE, LINEAR, EJ, false, false, true>(c, fa, st, err);
  }
  if (pg) {   // (one geometry per size: the peak-grouping kernels are the registers' other setting, not a tuning target)
    if constexpr (S == 14) return launch_fast_inst_lp<S, 5, CMODE, LINEAR, EJ, true>(c, fa, st, err);
    else return launch_fast_inst_lp<S, 4, CMODE, LINEAR, EJ, true>(c, fa, st, err);
  }
  if constexpr (S == 14) { if (fast_lp_runtime(S) == 5) return launch_fast_inst_lp<S, 5, CMODE, LINEAR, EJ>(c, fa, st, err); }
  return launch_fast_inst_lp<S, 4, CMODE, LINEAR, EJ>(c, fa, st, err);
}

template <int S, int EJ>
static rsp_status launch_fast_modes(int cmode, bool linear, bool pg, const LaunchCtx &c, const FastArgs &fa, cudaStream_t st, const char **err) {
  if (linear) switch (cmode) {
    case 0: return launch_fast_inst<S, 0, true, EJ>(c, fa, pg, st, err);
    case 1: return launch_fast_inst<S, 1, true, EJ>(c, fa, pg, st, err);
    default: return launch_fast_inst<S, 2, true, EJ>(c, fa, pg, st, err);
  }
  switch (cmode) {
    case 0: return launch_fast_inst<S, 0, false, EJ>(c, fa, pg, st, err);
    case 1: return launch_fast_inst<S, 1, false, EJ>(c, fa, pg, st, err);
    default: return launch_fast_inst<S, 2, false, EJ>(c, fa, pg, st, err);
  }
}

template <int S>
static rsp_status launch_fast_gos(const LaunchCtx &c, const FastArgs &fa, const GosArgs &ga, cudaStream_t st, const char **err) {
  using Gm = FGeo<S>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)Gm::CPC * GGeo<S>::ROW * sizeof(uint32_t);
  auto kernel = chain_fixed_fast_gos_kernel<S>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa, ga);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_fixed_fast_gos_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}

template <int S, int NWD, bool TW3, bool PG = false>
static rsp_status launch_fast_gosb_inst(const LaunchCtx &c, const FastArgs &fa, const GosArgs &ga, cudaStream_t st, const char **err) {
  using Gm = FGeo<S>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)Gm::CPC * (Gm::ROW + OGeo<S>::KW) * sizeof(uint32_t);
  auto kernel = chain_fixed_fast_gosb_kernel<S, NWD, TW3, PG>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa, ga);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_fixed_fast_gosb_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}
template <int S>
static rsp_status launch_fast_gosb(const LaunchCtx &c, const FastArgs &fa, const GosArgs &ga, cudaStream_t st, const char **err) {
  const int nwd = (fa.W + 31) / 32;
  const bool tw3 = fa.tw3 != nullptr;
  if (fa.pg_full >= 0) switch (nwd * 2 + (tw3 ? 1 : 0)) {
    case 2: return launch_fast_gosb_inst<S, 1, false, true>(c, fa, ga, st, err);
    case 3: return launch_fast_gosb_inst<S, 1, true, true>(c, fa, ga, st, err);
    case 4: return launch_fast_gosb_inst<S, 2, false, true>(c, fa, ga, st, err);
    case 5: return launch_fast_gosb_inst<S, 2, true, true>(c, fa, ga, st, err);
    case 6: return launch_fast_gosb_inst<S, 3, false, true>(c, fa, ga, st, err);
    case 7: return launch_fast_gosb_inst<S, 3, true, true>(c, fa, ga, st, err);
    case 8: return launch_fast_gosb_inst<S, 4, false, true>(c, fa, ga, st, err);
    default: return launch_fast_gosb_inst<S, 4, true, true>(c, fa, ga, st, err);
  }
  switch (nwd * 2 + (tw3 ? 1 : 0)) {
    case 2: return launch_fast_gosb_inst<S, 1, false>(c, fa, ga, st, err);
    case 3: return launch_fast_gosb_inst<S, 1, true>(c, fa, ga, st, err);
    case 4: return launch_fast_gosb_inst<S, 2, false>(c, fa, ga, st, err);
    case 5: return launch_fast_gosb_inst<S, 2, true>(c, fa, ga, st, err);
    case 6: return launch_fast_gosb_inst<S, 3, false>(c, fa, ga, st, err);
    case 7: return launch_fast_gosb_inst<S, 3, true>(c, fa, ga, st, err);
    case 8: return launch_fast_gosb_inst<S, 4, false>(c, fa, ga, st, err);
    default: return launch_fast_gosb_inst<S, 4, true>(c, fa, ga, st, err);
  }
}
// the rank-counting kernel needs W to be a multiple of 8 and >= 8; RSP_FAST_GOS_BITPLANES=0 selects it where it can run (A/B knob)
static bool gos_use_bitplanes(int W) {
  static int on = -1;
  if (on < 0) { const char *e = std::getenv("RSP_FAST_GOS_BITPLANES"); on = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return on || W < 8 || (W & 7);
}

template <int S, bool TW3>
static rsp_status launch_fast_os_inst(const LaunchCtx &c, const FastArgs &fa, const GosArgs &ga, cudaStream_t st, const char **err) {
  using Gm = FGeo<S>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)Gm::CPC * Gm::ROW * sizeof(uint32_t);
  auto kernel = chain_fixed_fast_os_kernel<S, TW3>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa, ga);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_fixed_fast_os_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}

template <int S>
static rsp_status launch_fast_os(const LaunchCtx &c, const FastArgs &fa, const GosArgs &ga, cudaStream_t st, const char **err) {
  return fa.tw3 != nullptr ? launch_fast_os_inst<S, true>(c, fa, ga, st, err) : launch_fast_os_inst<S, false>(c, fa, ga, st, err);
}

template <int S, int LP, bool TW3 = false>
static rsp_status launch_fast_fft_lp(const LaunchCtx &c, const FastArgs &fa, cudaStream_t st, const char **err) {
  using Gm = FGeo<S, LP>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)Gm::CPC * Gm::ROW * sizeof(uint32_t);
  auto kernel = fft_fixed_fast_kernel<S, LP, TW3>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa);
  if (cudaGetLastError() != cudaSuccess) { *err = "fft_fixed_fast_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}

// one transform size per compile unit
#define RSP_CAT2(a, b) a##b
#define RSP_CAT(a, b) RSP_CAT2(a, b)
rsp_status RSP_CAT(launch_fast_size_, RSP_FAST_S)(int core, int cmode, bool linear, int ej, const LaunchCtx &ctx, const FastArgs &fa, const GosArgs &ga,
                                                  cudaStream_t st, const char **err) {
  constexpr int S = RSP_FAST_S;
#if RSP_FAST_S < 10
  // 64 ... 512 points (the run-time sizes of a 1024-point generator): the CA family only, every bin on the edge path (a thread's 16 bins are
  // N / 16 apart: the window borders of most of them would need it anyway)
  if (core != CORE_MEM) { *err = "no OS / GOS variant below 1024 points"; return RSP_EUNSUPPORTED; }
  (void)ga; (void)ej;
  return launch_fast_modes<S, 8>(cmode, linear, fa.pg_full >= 0, ctx, fa, st, err);
#else
  if (core == CORE_LIS) return (gos_use_bitplanes(fa.W) || fa.tw3 != nullptr || fa.bpc || fa.cfar_only || fa.pg_full >= 0) ? launch_fast_gosb<S>(ctx, fa, ga, st, err) : launch_fast_gos<S>(ctx, fa, ga, st, err);
  if (core == CORE_OS) return launch_fast_os<S>(ctx, fa, ga, st, err);
#endif
#if RSP_FAST_S == 10
  if (ej == 2) return launch_fast_modes<S, 2>(cmode, linear, fa.pg_full >= 0, ctx, fa, st, err);
#endif
#if RSP_FAST_S >= 10
  return launch_fast_modes<S, 1>(cmode, linear, fa.pg_full >= 0, ctx, fa, st, err);
#endif
}

rsp_status RSP_CAT(launch_fast_fft_size_, RSP_FAST_S)(const LaunchCtx &ctx, const FastArgs &fa, cudaStream_t st, const char **err) {
  constexpr int S = RSP_FAST_S;
  if (fa.tw3 != nullptr) {
    if constexpr (S == 14) return launch_fast_fft_lp<S, 5, true>(ctx, fa, st, err);
    else return launch_fast_fft_lp<S, 4, true>(ctx, fa, st, err);
  }
  if constexpr (S == 14) { if (fast_lp_runtime(S) == 5) return launch_fast_fft_lp<S, 5>(ctx, fa, st, err); }
  return launch_fast_fft_lp<S, 4>(ctx, fa, st, err);
}

#if RSP_FAST_S == 10
#define RSP_DECL_FAST(s) rsp_status launch_fast_size_##s(int, int, bool, int, const LaunchCtx &, const FastArgs &, const GosArgs &, cudaStream_t, const char **); \
                         rsp_status launch_fast_fft_size_##s(const LaunchCtx &, const FastArgs &, cudaStream_t, const char **)
RSP_DECL_FAST(6); RSP_DECL_FAST(7); RSP_DECL_FAST(8); RSP_DECL_FAST(9); RSP_DECL_FAST(11); RSP_DECL_FAST(12); RSP_DECL_FAST(13); RSP_DECL_FAST(14);
#undef RSP_DECL_FAST

// the FFT block alone on the fast passes: the configuration of the fast chain's transform (16-bit data and twiddles, binPoint 0, RoundHalfUp,
// every stage scaled, forward, natural output order), any window
// constants of the binPoint >= 2 kernels (FastArgs::tw3 ...): the twiddle product's two trims, and for the chain the threshold's binary points
static void fast_fill_tw3(const FxArgs &a, FastArgs &fa) {
  const int bp_a = a.bp, bp_b = a.tw_bp, ctx = (bp_a > bp_b ? bp_a : bp_b) + 1;
  fa.tw3 = reinterpret_cast<const int2 *>(a.tww);
  fa.s1 = bp_a + bp_b - ctx; fa.s2 = ctx - bp_a;
  fa.h1 = 1 << (fa.s1 - 1); fa.h1m = fa.h1 - 1; fa.h2 = 1 << (fa.s2 - 1); fa.h2m = fa.h2 - 1;
}
// the CFAR threshold with binary points (rsp_device.cuh, cfar_bin): linear = the product trimmed (Floor) to the context's binary point
// max(inBP, scalerBP) + 1, log = the sum at max(inBP, scalerBP); the cell under test is compared at that binary point tbp and the threshold is
// reported at thresholdBP.  Returns false when a shift leaves the range the kernel's 32-bit arithmetic covers.
static bool fast_cfar_bp(const FxArgs &a, FastArgs *fa) {
  const int ib = a.in_bp, sb = a.sc_bp, tb = a.thr_bp;
  int thr_sh = 0, tbp;
  if (a.log_or_linear) { const int prod = ib + sb, ctx = (ib > sb ? ib : sb) + 1; thr_sh = prod > ctx ? prod - ctx : 0; tbp = prod - thr_sh; }
  else tbp = ib > sb ? ib : sb;
  const int cut_sh = tbp - ib, out_sh = tbp - tb, sc_sh = tbp - sb;
  if (ib < 0 || sb < 0 || tb < 0 || cut_sh < 0 || cut_sh > 14 || out_sh < 0 || out_sh > 30 || thr_sh > 30) return false;
  if (!a.log_or_linear) {
    if (sc_sh < 0 || sc_sh > 14) return false;
    const long long sp = (long long)a.scaler << sc_sh;
    if (sp >= (1LL << 30) || sp <= -(1LL << 30)) return false;
    if (fa) fa->sc_pre = (int)sp;
  } else if (fa) fa->sc_pre = a.scaler;
  if (fa) { fa->thr_sh = thr_sh; fa->ncut_mul = -(1 << cut_sh); fa->out_sh = out_sh; fa->t_mul = 1 << cut_sh; }
  return true;
}
static bool fast_tw3_ok(const Plan &pl, const FxArgs &a) {
  return pl.mode == MODE_FAST3 && a.bp >= 2 && a.bp <= 15 && a.tw_bp == 14 && a.tww != nullptr && !a.use4muls;
}

bool fast_fft_supported(const Plan &pl, const FxArgs &a) {
  if ((pl.mode != MODE_FAST && !fast_tw3_ok(pl, a)) || pl.first_stage != STAGE_FFT || pl.last_stage != STAGE_FFT) return false;
  if (pl.S < RSP_FAST_SMIN || pl.S > RSP_FAST_SMAX) return false;
  static int on = -1;
  if (on < 0) { const char *e = std::getenv("RSP_FAST_FFT"); on = (e != nullptr && e[0] == '0') ? 0 : 1; }   // A/B knob
  return on && (a.bp == 0 || pl.mode == MODE_FAST3) && a.tw_bp == 14 && a.bitrev && a.out != nullptr && a.in != nullptr;
}

rsp_status launch_fast_fft(const Plan &pl, const LaunchCtx &ctx, const FxArgs &a, cudaStream_t st, const char **err) {
  FastArgs fa = {};
  fa.in = a.in; fa.out = a.out; fa.n_rows = a.n_rows; fa.row_base = a.row_base;
  fa.tw = reinterpret_cast<const int2 *>(a.tw2);
  fa.win4 = a.win_on ? a.win4 : nullptr;
  for (int L = 0; L < 16; ++L) fa.tw_off[L] = a.tw_off[L];
  fa.one = 1; fa.m1 = -1; fa.rnd[0] = 0x8000; fa.rnd[1] = -0x8000; fa.rnd[2] = 0x7FFF; fa.rnd[3] = -0x7FFF;
  if (pl.mode == MODE_FAST3) fast_fill_tw3(a, fa);
  { static int l2p = -1; if (l2p < 0) { const char *e = std::getenv("RSP_FAST_L2_PREFETCH"); l2p = e ? std::atoi(e) : 1; }
    fa.l2_prefetch = (l2p != 0 && (reinterpret_cast<uintptr_t>(a.in) & 15u) == 0) ? 1 : 0; }
  switch (pl.S) {
  case 6: return launch_fast_fft_size_6(ctx, fa, st, err);
  case 7: return launch_fast_fft_size_7(ctx, fa, st, err);
  case 8: return launch_fast_fft_size_8(ctx, fa, st, err);
  case 9: return launch_fast_fft_size_9(ctx, fa, st, err);
  case 10: return launch_fast_fft_size_10(ctx, fa, st, err);
  case 11: return launch_fast_fft_size_11(ctx, fa, st, err);
  case 12: return launch_fast_fft_size_12(ctx, fa, st, err);
  case 13: return launch_fast_fft_size_13(ctx, fa, st, err);
  case 14: return launch_fast_fft_size_14(ctx, fa, st, err);
  default: *err = "no fast variant for this size"; return RSP_EUNSUPPORTED;
  }
}

static bool fast_cfar_bp(const FxArgs &a, FastArgs *fa);
static bool fast_bps_ok(const FxArgs &a) {   // binary points in the CFAR operands: the bit-plane kernels take them at run time
  if (a.in_bp == 0 && a.sc_bp == 0 && a.thr_bp == 0) return true;
  return a.cfar_trim == TRIM_FLOOR && fast_cfar_bp(a, nullptr);
}
static bool fast_gos_supported(const FxArgs &a) {
  if (a.core != CORE_LIS || a.cfar_algorithm != 1) return false;
  if (a.in_w != 16 || !fast_bps_ok(a) || a.send_cut) return false;   // (peak grouping: the <PG> instantiations)
  if (a.thr_w < 1 || a.thr_w > 32 || a.G < 0 || a.W < 1 || a.W > 128) return false;   // (W not a multiple of 8: bit planes only)
  if (a.idx_lagg < 1 || a.idx_lagg > a.W || a.idx_lead < 1 || a.idx_lead > a.W) return false;
  return a.out != nullptr && a.binw + 1 <= 31;
}

// the golden cfarOS on bit planes: up to 32 reference cells per side
static bool fast_os_supported(const FxArgs &a) {
  static int on = -1;
  if (on < 0) { const char *e = std::getenv("RSP_FAST_OS"); on = (e != nullptr && e[0] == '0') ? 0 : 1; }   // A/B knob
  if (!on || a.core != CORE_OS) return false;
  if (a.in_w != 16 || !fast_bps_ok(a) || a.send_cut || a.peak_grouping) return false;
  if (a.thr_w < 1 || a.thr_w > 32 || a.G < 0 || a.W < 1 || a.W > 32 || a.G + a.W > 64) return false;
  if (a.out == nullptr && a.det_list == nullptr) return false;
  return (reinterpret_cast<uintptr_t>(a.out) & 15u) == 0 && a.binw + 1 <= 31;
}

bool fast_chain_supported(const Plan &pl, const FxArgs &a) {
  // the CFAR block alone (rsp_cfar_run) with an ordered statistic: the bit-plane kernels on stored magnitude profiles
  if (pl.first_stage == STAGE_CFAR && pl.last_stage == STAGE_CFAR && pl.S >= 10 && pl.S <= RSP_FAST_SMAX && a.in != nullptr)
    return a.core == CORE_LIS ? fast_gos_supported(a) : (a.core == CORE_OS ? fast_os_supported(a) : false);
  const bool tw3 = fast_tw3_ok(pl, a);   // binPoint >= 2: the CA kernel's binary-point variant (FftMagCfarChainVanillaApp has 12)
  if ((pl.mode != MODE_FAST && !tw3) || pl.first_stage != STAGE_FFT || pl.last_stage != STAGE_CFAR) return false;
  if (pl.S < RSP_FAST_SMIN || pl.S > RSP_FAST_SMAX) return false;
  if ((a.bp != 0 && !tw3) || a.tw_bp != 14 || !a.bitrev || a.sel != 2 || a.mag_w != 16) return false;   // (the JPL magnitude does not look at the binary point)
  if (!tw3 && a.mag_bp != 0) return false;
  if (a.core == CORE_LIS) return pl.S >= 10 && fast_gos_supported(a);
  if (a.core == CORE_OS) return pl.S >= 10 && fast_os_supported(a);
  if (a.core != CORE_MEM || a.in_w != 16 || a.send_cut) return false;
  const bool bpc = tw3 || a.in_bp != 0 || a.sc_bp != 0 || a.thr_bp != 0;   // binary points in the CFAR: its own instantiation
  if (bpc && (a.cfar_trim != TRIM_FLOOR || !fast_cfar_bp(a, nullptr))) return false;   // (peak grouping with binary points: its own instantiations)
  if (a.out == nullptr && a.det_list == nullptr) return false;
  if (a.thr_w < 1 || a.thr_w > 32 || a.G < 1 || a.W < 1 || a.div_sum < 0 || a.div_sum > 30) return false;
  const int ej = fast_edge_iters(pl.S, a.W, a.G);
  if (pl.S >= 10 && ej > (pl.S == 10 ? 2 : 1)) return false;   // (below 1024 points every bin takes the edge path)
  // the scaled threshold must fit 32 bits (the window sums wrap at sumw bits like the reference's registers, see FastArgs::wmul)
  if (a.sumw < 2 || a.sumw > 31) return false;
  const long long smax = std::min<long long>((long long)a.W * 32768, 1LL << (a.sumw - 1));
  const long long tmax = (smax >> a.div_sum) + 1, sc = a.scaler < 0 ? -(long long)a.scaler : a.scaler;
  const long long lim = (1LL << 31) - (1LL << 16);   // the peak is the sign bit of thr - cut
  const int lsh = (a.in_bp > a.sc_bp ? a.in_bp : a.sc_bp);   // log mode adds at max(inBP, scalerBP)
  if (a.log_or_linear ? (tmax * sc >= lim) : ((tmax << (lsh - a.in_bp)) + (sc << (lsh - a.sc_bp)) >= lim)) return false;
  if (a.binw + 1 > 31) return false;
  return true;
}

rsp_status launch_fast_chain(const Plan &pl, const LaunchCtx &ctx, const FxArgs &a, cudaStream_t st, const char **err) {
  FastArgs fa;
  fa.in = a.in; fa.out = a.out; fa.n_rows = a.n_rows; fa.row_base = a.row_base;
  fa.tw = reinterpret_cast<const int2 *>(a.tw2);
  fa.win4 = a.win_on ? a.win4 : nullptr;
  for (int L = 0; L < 16; ++L) fa.tw_off[L] = a.tw_off[L];
  fa.W = a.W; fa.G = a.G; fa.scaler = a.scaler; fa.one = 1; fa.m1 = -1; fa.rnd[0] = 0x8000; fa.rnd[1] = -0x8000; fa.rnd[2] = 0x7FFF; fa.rnd[3] = -0x7FFF;
  fa.wmul = (int)(1u << (32 - a.sumw)); fa.div_sum = std::min(31, 32 - a.sumw + a.div_sum);
  fa.thr_mask = a.thr_w >= 32 ? 0xFFFFFFFFu : ((1u << a.thr_w) - 1u);
  fa.out_mul = (int)(1u << (a.binw + 1)); fa.thr_sext = 32 - a.thr_w;
  fa.det_list = a.det_list; fa.det_count = a.det_count; fa.det_cap = a.det_cap;
  fa.pg_full = a.peak_grouping ? (a.G == a.guard_window_size ? 1 : 0) : -1;   // -1: no peak grouping
  fa.cfar_only = pl.first_stage == STAGE_CFAR ? 1 : 0;
  fa.tw3 = nullptr; fa.bpc = 0; fa.thr_sh = 0; fa.ncut_mul = -1; fa.out_sh = 0; fa.t_mul = 1; fa.sc_pre = a.scaler;
  if (pl.mode == MODE_FAST3 && !fa.cfar_only) fast_fill_tw3(a, fa);
  if ((pl.mode == MODE_FAST3 && !fa.cfar_only) || a.in_bp != 0 || a.sc_bp != 0 || a.thr_bp != 0) { fast_cfar_bp(a, &fa); fa.bpc = 1; }
  // measured on B200 (profiles/r02_d_l2prefetch.txt): N = 16384 +5.3 %, N = 4096 +1.9 %, N = 1024 +1.3 %; RSP_FAST_L2_PREFETCH=0 turns it off
  { static int l2p = -1; if (l2p < 0) { const char *e = std::getenv("RSP_FAST_L2_PREFETCH"); l2p = e ? std::atoi(e) : 1; }
    fa.l2_prefetch = (l2p != 0 && a.gen_full == nullptr && (reinterpret_cast<uintptr_t>(a.in) & 15u) == 0) ? (l2p > 4 ? 4 : l2p) : 0; }   // value = rows ahead
  fa.gen_phase = a.gen_phase; fa.gen_full = a.core == CORE_MEM ? a.gen_full : nullptr; fa.gen_period = a.gen_period; fa.gen_pp = a.gen_pp;
  fa.gen_mask = a.gen_mask; fa.gen_k0 = a.gen_k0;
  { static int skew = -1; if (skew < 0) { const char *e = std::getenv("RSP_FAST_SKEW_NS"); skew = e ? std::atoi(e) : 0; } fa.skew_ns = skew; }
  fa.stage = (pl.S < RSP_FAST_TWSMEM_S && fast_stage_enabled(pl.S) && a.gen_full == nullptr && (reinterpret_cast<uintptr_t>(a.in) & 15u) == 0) ? 1 : 0;   // rows are N * 4 bytes: the base decides
  const int cmode = a.cfar_mode == 0 ? 0 : (a.cfar_mode == 1 ? 1 : 2);   // MuxLookup default: smallestOf (CFARCoreWithMem.scala:138-141)
  const bool linear = a.log_or_linear != 0;
  const GosArgs ga = {cmode, linear ? 1 : 0, a.idx_lagg - 1, a.idx_lead - 1};
  const int ej = a.core != CORE_MEM ? 1 : fast_edge_iters(pl.S, a.W, a.G);
  switch (pl.S) {
  case 6: return launch_fast_size_6(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 7: return launch_fast_size_7(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 8: return launch_fast_size_8(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 9: return launch_fast_size_9(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 10: return launch_fast_size_10(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 11: return launch_fast_size_11(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 12: return launch_fast_size_12(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 13: return launch_fast_size_13(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  case 14: return launch_fast_size_14(a.core, cmode, linear, ej, ctx, fa, ga, st, err);
  default: *err = "no fast variant for this size"; return RSP_EUNSUPPORTED;
  }
}
#endif

}  // namespace rsp
