// rsp_device.cuh -- per-thread device code of the rsp-chains hot path (FFT -> LogMagMux -> CFAR).
//
// Every function here is the work of ONE thread during ONE phase of a chirp; phases are separated by a
// block-wide barrier.  The CUDA kernels (rsp_kernels.cuh) call the phases with threadIdx-derived indices.  The functions are
// host-callable as well (RSP_HD): the host builds tables with the very code the device reads them with (nco_lookup), and
// tests/emu/emu.cpp (test infrastructure) compiles this header with g++ and runs the phases thread by thread on argument blocks from
// the product's planner, so index maps, integer arithmetic and planning are held to the oracle on a box without a GPU
// (tests/test_emu_cpu.py).  The product library only ever runs the CUDA kernels.
//
// Arithmetic follows the reference generators (paths relative to the rsp-chains repository):
//   butterfly + div2        generators/sdf-fft/src/main/scala/SDFFFTUtil.scala:51-61, SDFChainRadix22.scala:459-464,487-493
//   -j / twiddle multiply   generators/sdf-fft/src/main/scala/SDFChainRadix22.scala:224-352
//   bit reversal            generators/sdf-fft/src/main/scala/BitReversePingPong.scala:93-109
//   magnitude / log2        generators/logMagMux/src/main/scala/LogMagMux.scala:92-169
//   CFAR cores              generators/cfar/src/main/scala/CFARCoreWithMem.scala:102-189, CFARCoreWithLis.scala:136-194,
//                           CFARCoreWithASR.scala:193-209,285-332; beat packing CFARDspBlock.scala:49,78-90
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RSP_HD __host__ __device__ __forceinline__
#else
#define RSP_HD inline
#endif

namespace rsp {

enum { TRIM_ROUND_HALF_UP = 0, TRIM_FLOOR = 1, TRIM_CONVERGENT = 2, TRIM_ROUND_UP = 3, TRIM_ROUND_TOWARDS_ZERO = 4,
       TRIM_ROUND_TOWARDS_INFINITY = 5, TRIM_ROUND_HALF_DOWN = 6, TRIM_ROUND_HALF_TOWARDS_ZERO = 7,
       TRIM_ROUND_HALF_TOWARDS_INFINITY = 8, TRIM_ROUND_HALF_TO_ODD = 9, TRIM_NO_TRIM = 10 };
enum { CORE_MEM = 0, CORE_LIS = 1, CORE_ASR = 2,
       CORE_OS = 3 };   // the combined-window ordered statistic of the Scala golden cfarOS (CFARUtils.scala:159-196); no hardware core
enum { STAGE_FFT = 0, STAGE_MAG = 1, STAGE_CFAR = 2 };
// kernel variants of the fixed-point chain
enum { MODE_GENERIC = 0,   // every parameter read at run time
       MODE_FAST = 1,      // 16-bit data, RoundHalfUp, scale every stage, forward, product trimmed once (binPoint <= 1)
       MODE_FAST3 = 2,     // as FAST but 3-multiplier product with per-product trim (binPoint >= 2, use4Muls = false)
       MODE_RADIX2 = 3,    // sdfRadix "2" (SDFChainRadix2.scala), DIF or DIT; every parameter read at run time like GENERIC
       MODE_FASTC = 4,     // as FAST with trimType = Convergent (what the reference's SQNR specs elaborate), rounding folded at compile time
       MODE_FAST3C = 5 };  // as FAST3 with trimType = Convergent
constexpr bool mode_is_generic(int m) { return m == MODE_GENERIC || m == MODE_RADIX2; }
constexpr bool mode_is_convergent(int m) { return m == MODE_FASTC || m == MODE_FAST3C; }

struct Detection { uint32_t chirp, bin; int32_t threshold, cut; };

// arguments shared by all fixed-point kernels
struct FxArgs {
  const uint32_t *in;
  uint32_t *out;
  uint32_t n_rows;          // chirps in this launch
  uint32_t row_base;        // index of row 0 inside the caller's batch (reported in detections)
  int first_stage, last_stage;
  // FFT
  int data_width, bp, tw_bp, trim, use4muls, bitrev, fft_dir;   // bitrev: the result is stored at the bit-reversed index
  int radix2, dit, load_rev;   // sdfRadix "2" (SDFChainRadix2.scala), DITDecimType, beats arrive in bit-reversed order (DIT without useBitReverse)
  int dit_tw_par;              // radix-2^2 DIT: stage i > 0 carries the twiddle multiplier iff (i + dit_tw_par) is even, else the x(-j)
  uint32_t keep_mask;       // bit i: stage i of THIS transform scales (1) or keeps LSBs (0)
  const uint32_t *tw;       // packed (cos16 << 16 | sin16) twiddles, per stageN three runs [W^2k | W^k | W^3k]
  int tw_off[16];           // word offset of the table of stageN = 1 << L
  const int32_t *tw2;       // same entries as int32 pairs (c, s) << (16 - tw_bp) (fast kernels, rsp_fast.cu)
  const int32_t *tww;       // same entries as raw int32 pairs (fast kernels with binPoint >= 2; null when not uploaded)
  const int32_t *win;       // window ROM (protoWin raw values), read when win_on
  const int32_t *win4;      // the same << (16 - win_shift): bits 16..31 of beat * win4 are the windowed beat (fast kernels)
  int win_on, win_shift;    // win_shift = binary point of protoWin = windowWidth - 2
  // LogMagMux
  int sel, mag_w, mag_bp, log_w, log_bp, log_L, mag_trim;
  const int32_t *log_tab;
  // CFAR
  int core, in_w, in_bp, thr_w, thr_bp, sc_w, sc_bp, sumw, guard_window_size, binw, send_cut;
  int scaler, log_or_linear, div_sum, peak_grouping, cfar_algorithm, cfar_mode, W, G, idx_lagg, idx_lead;
  int sub_cells, sub_log2, cfar_trim;
  Detection *det_list;
  uint32_t *det_count;
  uint32_t det_cap;
  // per-stage overflow status (FFTParams.overflowReg, SDFChainRadix22.scala:473-486): one device word, bit (ovf_shift + i) is OR-ed
  // in when a butterfly of stage i of this transform has a grown output at or above 2^(dataWidth - 1); null = not elaborated
  uint32_t *ovf_flags;
  int ovf_shift;
  // PLFG -> NCO fused into the first pass of the fast chain kernel (rsp_fast.cu, FastArgs::gen_*); gen_full null = beats come from `in`
  const uint16_t *gen_phase;
  const uint32_t *gen_full;
  uint32_t gen_period, gen_pp, gen_mask;
  unsigned long long gen_k0;
};

// ------------------------------------------------------------------------------------------------ helpers
RSP_HD int wrapw(int v, int w) { const int s = 32 - w; return (int)((uint32_t)v << s) >> s; }
RSP_HD long long wrapw64(long long v, int w) { const int s = 64 - w; return (long long)((unsigned long long)v << s) >> s; }
RSP_HD int wrap16(int v) { return (int)(short)v; }
RSP_HD uint32_t pack_iq(int re, int im) { return ((uint32_t)re << 16) | ((uint32_t)im & 0xFFFFu); }
RSP_HD int hi16(uint32_t w) { return (int)w >> 16; }
RSP_HD int lo16(uint32_t w) { return (int)(short)(w & 0xFFFFu); }
RSP_HD int pad(int p) { return p + (p >> 5); }

RSP_HD uint32_t bitrev_n(uint32_t v, int bits) {
#if defined(__CUDA_ARCH__)
  return __brev(v) >> (32 - bits);
#else
  uint32_t r = 0;
  for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1u) << (bits - 1 - i);
  return r;
#endif
}
RSP_HD int ilog2_floor_u64(unsigned long long v) {
#if defined(__CUDA_ARCH__)
  return v ? 63 - __clzll((long long)v) : 0;
#else
  int n = 0; while (v >>= 1) ++n; return n;
#endif
}
RSP_HD int bitlen_u32(uint32_t v) { int n = 0; while (v) { ++n; v >>= 1; } return n; }
RSP_HD int ilog2_ceil_u32(uint32_t v) { int n = 0; while ((1u << n) < v) ++n; return n; }

template <typename T> RSP_HD T ldg(const T *p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

// v / 2^shift under a dsptools TrimType (div2 / trimBinary)
template <typename I> RSP_HD I trim_rt(I v, int shift, int mode) {
  if (shift <= 0 || mode == TRIM_NO_TRIM) return v;
  const I one = 1, half = one << (shift - 1), mask = (one << shift) - 1;
  const I q = v >> shift, r = v & mask;
  switch (mode) {
  case TRIM_ROUND_HALF_UP: return (v + half) >> shift;
  case TRIM_FLOOR: return q;
  case TRIM_CONVERGENT: return r > half ? q + 1 : (r == half ? q + (q & 1) : q);
  case TRIM_ROUND_UP: return (v + mask) >> shift;
  case TRIM_ROUND_TOWARDS_ZERO: return v >= 0 ? q : -((-v) >> shift);
  case TRIM_ROUND_TOWARDS_INFINITY: return v >= 0 ? ((v + mask) >> shift) : -((-v + mask) >> shift);
  case TRIM_ROUND_HALF_DOWN: return (v + half - 1) >> shift;
  case TRIM_ROUND_HALF_TOWARDS_ZERO: return v >= 0 ? ((v + half - 1) >> shift) : -((-v + half - 1) >> shift);
  case TRIM_ROUND_HALF_TOWARDS_INFINITY: return v >= 0 ? ((v + half) >> shift) : -((-v + half) >> shift);
  case TRIM_ROUND_HALF_TO_ODD: return r > half ? q + 1 : (r == half ? q + (1 - (q & 1)) : q);
  default: return q;
  }
}

// Convergent (round half to even) with the shift known to be >= 1: the compile-time form of trim_rt's TRIM_CONVERGENT case
RSP_HD int trim_conv(int v, int shift) {
  const int half = 1 << (shift - 1), q = v >> shift, r = v & ((1 << shift) - 1);
  return q + ((r > half || (r == half && (q & 1))) ? 1 : 0);
}
RSP_HD int trim_conv1(int v) { return (v >> 1) + ((v & (v >> 1)) & 1); }   // shift 1: the remainder is the half exactly when v is odd

// ------------------------------------------------------------------------------------------------ geometry
constexpr int fx_n_passes(int S) { return S <= 4 ? 1 : ((S % 4 == 0) ? S / 4 : S / 4 + 1); }
constexpr int fx_pass_size(int S, int k) {
  return S <= 4 ? S
       : (S % 4 == 0) ? 4
       : (S % 4 == 2) ? (k < S / 4 ? 4 : 2)
       : (S % 4 == 3) ? (k < S / 4 ? 4 : 3)
       : (k < S / 4 - 1 ? 4 : (k == S / 4 - 1 ? 2 : 3));
}
constexpr int fx_pass_start(int S, int k) { return k == 0 ? 0 : fx_pass_start(S, k - 1) + fx_pass_size(S, k - 1); }

template <int S> struct Geo {
  static constexpr int N = 1 << S;
  static constexpr int PTS = N < 16 ? N : 16;          // points per thread
  static constexpr int TPC = N / PTS;                  // threads per chirp
  static constexpr int CPC = TPC >= 256 ? 1 : 256 / TPC; // chirps per CTA
  static constexpr int THREADS = TPC * CPC;
  static constexpr int BUF = N + (N >> 5) + 2;         // padded words of one chirp buffer (N + 1 entries for the prefix sums)
  static constexpr int GRP = TPC < 32 ? TPC : 32;      // scan group width
  static constexpr int NGRP = TPC / GRP;               // scan groups per chirp
};

// ------------------------------------------------------------------------------------------------ FFT arithmetic
// returns `tooBig` of the reference's overflow check on the four grown outputs (generic modes with the status register elaborated);
// `tooSmall` can never fire: an arithmetic right shift of a negative value is never 0 (SDFChainRadix22.scala:476-479)
template <int MODE> RSP_HD bool butterfly(const FxArgs &a, int scale, int &ar, int &ai, int &cr, int &ci) {
  const int sr = ar + cr, si = ai + ci, dr = ar - cr, di = ai - ci;
  bool ov = false;
  if (mode_is_generic(MODE) && a.ovf_flags != nullptr) {
    const int lim = 1 << (a.data_width - 1);
    ov = sr >= lim || si >= lim || dr >= lim || di >= lim;
  }
  if (mode_is_convergent(MODE)) {
    ar = trim_conv1(sr); ai = trim_conv1(si);
    cr = wrap16(trim_conv1(dr)); ci = wrap16(trim_conv1(di));
  } else if (!mode_is_generic(MODE)) {
    // RoundHalfUp div2: (v + 1) >> 1.  The sum cannot leave 16 bits; the difference can (+32768) and wraps.
    ar = (sr + 1) >> 1; ai = (si + 1) >> 1;
    cr = wrap16((dr + 1) >> 1); ci = wrap16((di + 1) >> 1);
  } else {
    const int w = a.data_width;
    if (scale) {
      ar = wrapw(trim_rt<int>(sr, 1, a.trim), w); ai = wrapw(trim_rt<int>(si, 1, a.trim), w);
      cr = wrapw(trim_rt<int>(dr, 1, a.trim), w); ci = wrapw(trim_rt<int>(di, 1, a.trim), w);
    } else {
      ar = wrapw(sr, w); ai = wrapw(si, w); cr = wrapw(dr, w); ci = wrapw(di, w);
    }
  }
  return ov;
}

// multiply by -j: (re, im) <- (im, -re)
template <int MODE> RSP_HD void mul_mj(const FxArgs &a, int &re, int &im) {
  const int r = re;
  re = im;
  im = (!mode_is_generic(MODE)) ? wrap16(-r) : wrapw(-r, a.data_width);
}

// DspComplex context_* then trimBinary(bpos) then `:=` into protoIQ
template <int MODE> RSP_HD void cmul(const FxArgs &a, int &xr, int &xi, uint32_t tw) {
  const int br = hi16(tw), bi = lo16(tw);
  const int ar = xr, ai = xi;
  if (MODE == MODE_FAST) {
    // product binary point bp + tw_bp <= context binary point: nothing is trimmed before the final trimBinary,
    // so 3- and 4-multiplier forms give the same exact sums
    const int sh = a.tw_bp, half = 1 << (sh - 1);
    xr = wrap16((ar * br - ai * bi + half) >> sh);
    xi = wrap16((ar * bi + ai * br + half) >> sh);
  } else if (MODE == MODE_FASTC) {
    const int sh = a.tw_bp;
    xr = wrap16(trim_conv(ar * br - ai * bi, sh));
    xi = wrap16(trim_conv(ar * bi + ai * br, sh));
  } else if (MODE == MODE_FAST3C) {
    const int bp_a = a.bp, bp_b = a.tw_bp;
    const int ctx_bp = (bp_a > bp_b ? bp_a : bp_b) + 1;
    const int s1 = bp_a + bp_b - ctx_bp, s2 = ctx_bp - bp_a;
    const int k1 = trim_conv(ar * (br + bi), s1);
    const int k2 = trim_conv((ar + ai) * bi, s1);
    const int k3 = trim_conv((ai - ar) * br, s1);
    xr = wrap16(trim_conv(k1 - k2, s2));
    xi = wrap16(trim_conv(k1 + k3, s2));
  } else if (MODE == MODE_FAST3) {
    const int bp_a = a.bp, bp_b = a.tw_bp;
    const int ctx_bp = (bp_a > bp_b ? bp_a : bp_b) + 1;
    const int s1 = bp_a + bp_b - ctx_bp, h1 = 1 << (s1 - 1);
    const int s2 = ctx_bp - bp_a, h2 = 1 << (s2 - 1);
    const int k1 = (ar * (br + bi) + h1) >> s1;
    const int k2 = ((ar + ai) * bi + h1) >> s1;
    const int k3 = ((ai - ar) * br + h1) >> s1;
    xr = wrap16((k1 - k2 + h2) >> s2);
    xi = wrap16((k1 + k3 + h2) >> s2);
  } else {
    const int bp_a = a.bp, bp_b = a.tw_bp, trim = a.trim;
    const int prod_bp = bp_a + bp_b;
    const int ctx_bp = (bp_a > bp_b ? bp_a : bp_b) + 1;
    const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
    const int cur_bp = prod_bp - s1;
    int re, im;
    if (a.use4muls) {
      re = trim_rt<int>(ar * br, s1, trim) - trim_rt<int>(ai * bi, s1, trim);
      im = trim_rt<int>(ar * bi, s1, trim) + trim_rt<int>(ai * br, s1, trim);
    } else {
      const int k1 = trim_rt<int>(ar * (br + bi), s1, trim);
      const int k2 = trim_rt<int>((ar + ai) * bi, s1, trim);
      const int k3 = trim_rt<int>((ai - ar) * br, s1, trim);
      re = k1 - k2;
      im = k1 + k3;
    }
    if (cur_bp > bp_a) { re = trim_rt<int>(re, cur_bp - bp_a, trim); im = trim_rt<int>(im, cur_bp - bp_a, trim); }
    else if (cur_bp < bp_a) { re <<= (bp_a - cur_bp); im <<= (bp_a - cur_bp); }
    xr = wrapw(re, a.data_width);
    xi = wrapw(im, a.data_width);
  }
}

// radix-2 chain: the product is exact (trimType = NoTrim, overflowType = Grow) and trimmed once by trimBinary(bpos) under
// params.trimType, then assigned to protoIQ (SDFChainRadix2.scala:327-330,349-352)
RSP_HD void cmul_r2(const FxArgs &a, int &xr, int &xi, uint32_t tw) {
  const int br = hi16(tw), bi = lo16(tw);
  const int ar = xr, ai = xi;
  xr = wrapw(trim_rt<int>(ar * br - ai * bi, a.tw_bp, a.trim), a.data_width);
  xi = wrapw(trim_rt<int>(ar * bi + ai * br, a.tw_bp, a.trim), a.data_width);
}

// The MODE_RADIX2 kernels (three stage structures x every trim type, unrolled 16 points deep) call the two products out of line:
// they are parity paths, and inlining both products into every unrolled point made their compile units the longest of the build.
#if defined(__CUDACC__)
static __device__ __noinline__ int2 cmul_r2_call(int xr, int xi, uint32_t tw, int tw_bp, int trim, int data_width) {
  FxArgs a;
  a.tw_bp = tw_bp; a.trim = trim; a.data_width = data_width;
  cmul_r2(a, xr, xi, tw);
  return make_int2(xr, xi);
}
static __device__ __noinline__ int2 cmul_generic_call(int xr, int xi, uint32_t tw, int bp, int tw_bp, int trim, int use4muls, int data_width) {
  FxArgs a;
  a.bp = bp; a.tw_bp = tw_bp; a.trim = trim; a.use4muls = use4muls; a.data_width = data_width;
  cmul<MODE_GENERIC>(a, xr, xi, tw);
  return make_int2(xr, xi);
}
#define RSP_CMUL_R2(a, re, im, tw) do { const int2 r_ = cmul_r2_call(re, im, tw, (a).tw_bp, (a).trim, (a).data_width); re = r_.x; im = r_.y; } while (0)
#define RSP_CMUL_GEN(a, re, im, tw) do { const int2 r_ = cmul_generic_call(re, im, tw, (a).bp, (a).tw_bp, (a).trim, (a).use4muls, (a).data_width); re = r_.x; im = r_.y; } while (0)
#else
#define RSP_CMUL_R2(a, re, im, tw) cmul_r2(a, re, im, tw)
#define RSP_CMUL_GEN(a, re, im, tw) cmul<MODE_GENERIC>(a, re, im, tw)
#endif

// ------------------------------------------------------------------------------------------------ LogMagMux
// returns the low half-word value (.imag of the output beat); *out_re is the high half (.real)
RSP_HD int mag_one(const FxArgs &a, int re, int im, int *out_re) {
  const int w = a.mag_w, bp = a.mag_bp;
  *out_re = 0;
  if (a.sel == 0) { *out_re = re; return im; }
  const long long ar = re < 0 ? -(long long)re : re, ai = im < 0 ? -(long long)im : im;
  if (a.sel == 1) {
    const int prod_bp = 2 * bp, ctx_bp = bp + 1;
    const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
    const int cur_bp = prod_bp - s1;
    long long v = trim_rt<long long>(ar * ar, s1, a.mag_trim) + trim_rt<long long>(ai * ai, s1, a.mag_trim);
    if (cur_bp > bp) v >>= (cur_bp - bp); else if (cur_bp < bp) v <<= (bp - cur_bp);
    return (int)wrapw64(v, w);
  }
  const int u = (int)(ar > ai ? ar : ai), v = (int)(ar > ai ? ai : ar);
  const int op1 = u + (v >> 3);
  const int op2 = (u - (u >> 3)) + (v >> 1);
  const int jpl = op1 > op2 ? op1 : op2;
  if (a.sel == 2) return wrapw(jpl, w);
  // sel 3: log2 of the JPL magnitude, per-sample arithmetic (see DESIGN.md for the RTL pipeline skew)
  const int wj = w + 2;
  const unsigned long long ju = (unsigned long long)jpl & ((1ULL << wj) - 1);
  const int k = ilog2_floor_u64(ju);
  const int wk_a = ilog2_ceil_u32((uint32_t)wj);
  int wk = wk_a;
  { int wb = bitlen_u32((uint32_t)bp); if (wb < 1) wb = 1; if (wb > wk) wk = wb; }
  const long long log_sint = wrapw64((long long)((uint32_t)(k - bp) & ((1u << wk) - 1)), wk);
  const unsigned long long no_lead_one = (ju - (1ULL << k)) & ((1ULL << wj) - 1);
  int ws = bitlen_u32((uint32_t)w); if (wk_a > ws) ws = wk_a;
  const uint32_t shift_num = (uint32_t)(w - k - 1) & ((1u << ws) - 1);
  const int L = a.log_L;
  const unsigned long long shifted = shift_num < 63 ? (no_lead_one << shift_num) : 0;
  const uint32_t addr = (uint32_t)((shifted >> (w - L - 1)) & ((1ULL << L) - 1));
  const long long frac = ldg(&a.log_tab[addr]);
  const long long lg = wrapw64((log_sint << a.log_bp) + frac, a.log_w);
  const unsigned long long raw = (unsigned long long)lg & ((1ULL << a.log_w) - 1);
  return (int)(a.log_w >= w ? wrapw64((long long)raw, w) : (long long)raw);
}

// ------------------------------------------------------------------------------------------------ FFT passes
// One pass = m consecutive radix-2 stages done in registers on 2^m points that are Dl apart.  Passes start on an
// even stage index so every radix-2^2 pair (butterfly, -j, butterfly, twiddle) stays inside one pass.
// gin: the chirp in global memory (pass 0 only); bufA: padded complex words; bufM: natural-order result beats.
template <int S, int MODE, int K>
RSP_HD void fft_pass(const FxArgs &a, int tl, const uint32_t *gin, uint32_t *bufA, uint32_t *bufM) {
  constexpr int m = fx_pass_size(S, K), i0 = fx_pass_start(S, K), LDl = S - i0 - m, Dl = 1 << LDl;
  constexpr int PTS = Geo<S>::PTS, G = PTS >> m, R = 1 << m;
  constexpr bool first = (K == 0), last = (K == fx_n_passes(S) - 1);
  static_assert((i0 & 1) == 0, "passes start on even stages");
  static_assert(!last || LDl == 0, "last pass ends the transform");
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int vt = tl * G + g;
    const int low = vt & (Dl - 1), high = vt >> LDl;
    const int base = (high << (m + LDl)) | low;
    int re[R], im[R];
    uint32_t ovf = 0;
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int p = base + rr * Dl;
      const int gp = (mode_is_generic(MODE) && first && a.load_rev) ? (int)bitrev_n((uint32_t)p, S) : p;
      const uint32_t wd = first ? ldg(&gin[gp]) : bufA[pad(p)];
      if (!mode_is_generic(MODE)) {   // dataWidth 16, forward
        re[rr] = hi16(wd); im[rr] = lo16(wd);
        if (first && a.win_on) {
          const int w = ldg(&a.win[p]);
          re[rr] = wrap16((re[rr] * w) >> a.win_shift); im[rr] = wrap16((im[rr] * w) >> a.win_shift);
        }
      } else {
        int r0 = wrapw((int)(wd >> 16), a.data_width), i0v = wrapw((int)(wd & 0xFFFFu), a.data_width);
        if (first && a.win_on) {
          // in * window(cntWin): full-precision product, then the assignment to protoIQ drops win_shift fractional bits
          // (floor) and the MSBs beyond dataWidth (SDFFFT.scala:233-234)
          const int w = ldg(&a.win[p]);
          r0 = wrapw((r0 * w) >> a.win_shift, a.data_width); i0v = wrapw((i0v * w) >> a.win_shift, a.data_width);
        }
        // ifft swaps real and imaginary on the way in (SDFChainRadix22.scala:205-211)
        const bool sw = first && !a.fft_dir;
        re[rr] = sw ? i0v : r0; im[rr] = sw ? r0 : i0v;
      }
    }
#pragma unroll
    for (int s = 0; s < m; ++s) {
      const int i = i0 + s;
      const int dist = 1 << (m - 1 - s);
      const int scale = (!mode_is_generic(MODE)) ? 1 : (int)((a.keep_mask >> i) & 1u);
      if (MODE == MODE_RADIX2) {
        // Radix-2 SDF chain.  DIF: the stage of delay D = Dl * dist multiplies the difference outputs k = 1..D-1 of every block by
        // W_2D^k behind the butterfly (SDFChainRadix2.scala:343-357).  DIT is run in its natural-in / bit-reversed-out form, the
        // same index space as DIF (point q here is point bitrev(q) of the reference's stream): stage i pairs points 2^(S-1-i)
        // apart and multiplies the second one by W_2^(i+1)^k IN FRONT of the butterfly (:321-339), k = bitrev_i(block number).
        // Both ROMs are run 0 ([W^2k], k < stageN/4) of the table of stageN = 4 * delay.
        const uint32_t *tab = a.tw + a.tw_off[a.dit ? i + 2 : S - i + 1];
        if (a.dit && i > 0 && a.radix2) {
#pragma unroll
          for (int rr = 0; rr < R; ++rr)
            if (rr & dist) {
              const int blk = (high << s) | (rr >> (m - s));
              const int k = (int)bitrev_n((uint32_t)blk, i);
              if (k) RSP_CMUL_R2(a, re[rr], im[rr], ldg(&tab[k]));
            }
        } else if (a.dit && i > 0) {
          // Radix-2^2 DIT (SDFChainRadix22.scala:224-352), same index space: bit b of the reference's position p is bit S-1-b of q.
          // Twiddle stages multiply EVERY input by [0]*D ++ [2k] ++ [k] ++ [3k] over p mod 4D (quarter = bits i, i+1 of p, k = the
          // i bits below), the others turn positions [3D/2, 2D) of each 2D block (bits i and i-1 of p set) by -j.
          const bool is_tw = ((i + a.dit_tw_par) & 1) == 0;
#pragma unroll
          for (int rr = 0; rr < R; ++rr) {
            const int qi = base + rr * Dl;
            const int b0 = (rr & dist) ? 1 : 0;
            if (is_tw) {
              const int b1 = i == S - 1 ? 0 : ((qi >> (S - 2 - i)) & 1);
              const int quarter = b0 + 2 * b1;
              if (quarter) {
                const int k = (int)bitrev_n((uint32_t)(qi >> (S - i)), i);
                const int slot = quarter == 1 ? 0 : (quarter == 2 ? 1 : 2);
                RSP_CMUL_GEN(a, re[rr], im[rr], ldg(&tab[(slot << i) + k]));
              }
            } else if (b0 && ((qi >> (S - i)) & 1)) {
              mul_mj<MODE>(a, re[rr], im[rr]);
            }
          }
        }
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
          if (!(rr & dist)) { if (butterfly<MODE>(a, scale, re[rr], im[rr], re[rr + dist], im[rr + dist])) ovf |= 1u << i; }
        if (!a.dit && i != S - 1) {
#pragma unroll
          for (int rr = 0; rr < R; ++rr)
            if (rr & dist) {
              const int k = (rr & (dist - 1)) * Dl + low;
              if (k) RSP_CMUL_R2(a, re[rr], im[rr], ldg(&tab[k]));
            }
        }
        continue;
      }
#pragma unroll
      for (int rr = 0; rr < R; ++rr)
        if (!(rr & dist)) { if (butterfly<MODE>(a, scale, re[rr], im[rr], re[rr + dist], im[rr + dist])) ovf |= 1u << i; }
      if (i == S - 1) continue;
      if ((i & 1) == 0) {
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
          if ((rr & dist) && (rr & (dist >> 1))) mul_mj<MODE>(a, re[rr], im[rr]);
      } else {
        const int L = S - i + 1;                 // stageN = 1 << L = 4 D
        const int q = 1 << (L - 2);              // = D = Dl * dist
        const uint32_t *tab = a.tw + a.tw_off[L];
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
          const int quarter = ((rr >> (m - 1 - s)) & 1) + 2 * ((rr >> (m - s)) & 1);
          if (quarter) {
            const int k = (rr & (dist - 1)) * Dl + low;
            const int slot = quarter == 1 ? 0 : (quarter == 2 ? 1 : 2);   // [W^2k | W^k | W^3k]
            cmul<MODE>(a, re[rr], im[rr], ldg(&tab[slot * q + k]));
          }
        }
      }
    }
    if (mode_is_generic(MODE) && ovf != 0) {
#if defined(__CUDA_ARCH__)
      atomicOr(a.ovf_flags, ovf << a.ovf_shift);
#else
      *a.ovf_flags |= ovf << a.ovf_shift;
#endif
    }
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int p = base + rr * Dl;
      if (!last) { bufA[pad(p)] = pack_iq(re[rr], im[rr]); continue; }
      const int bin = a.bitrev ? (int)bitrev_n((uint32_t)p, S) : p;
      int orr = re[rr], oi = im[rr];
      if (mode_is_generic(MODE) && !a.fft_dir) { const int t = orr; orr = oi; oi = t; }
      uint32_t wd = pack_iq(orr, oi);
      if (a.last_stage != STAGE_FFT) {
        int mr, mi, ore;
        if (!mode_is_generic(MODE)) { mr = orr; mi = oi; }
        else { mr = wrapw((int)(wd >> 16), a.mag_w); mi = wrapw((int)(wd & 0xFFFFu), a.mag_w); }
        const int v = mag_one(a, mr, mi, &ore);
        wd = pack_iq(ore, v);
      }
      bufM[pad(bin)] = wd;
    }
  }
}

// ------------------------------------------------------------------------------------------------ CFAR phases
// x[n] = low in_w bits of the beat (CFARDspBlock.scala:49)
RSP_HD int cfar_x(const FxArgs &a, const uint32_t *bufM, int n) { return wrapw((int)bufM[pad(n)], a.in_w); }

// phase A: each thread sums its PTS consecutive cells
template <int S> RSP_HD int cfar_thread_total(const FxArgs &a, int tl, const uint32_t *bufM) {
  constexpr int PTS = Geo<S>::PTS;
  int t = 0;
#pragma unroll
  for (int j = 0; j < PTS; ++j) t += cfar_x(a, bufM, tl * PTS + j);
  return t;
}
// phase C: write the exclusive prefix sums P[n] of the thread's cells (and P[N] by the last thread)
template <int S> RSP_HD void cfar_write_prefix(const FxArgs &a, int tl, int offset, const uint32_t *bufM, uint32_t *bufP) {
  constexpr int PTS = Geo<S>::PTS, N = Geo<S>::N;
  int acc = offset;
#pragma unroll
  for (int j = 0; j < PTS; ++j) {
    const int n = tl * PTS + j;
    bufP[pad(n)] = (uint32_t)acc;
    acc += cfar_x(a, bufM, n);
  }
  if (tl == Geo<S>::TPC - 1) bufP[pad(N)] = (uint32_t)acc;
}

// k-th largest (1-based) of W cells starting at `start`: the value the descending linear insertion sorter holds at
// index k-1 (CFARCoreWithLis.scala:32-39,136-143).  Bitwise search on the biased value.
// `start2` >= 0: the window is the union of [start, start + W) and [start2, start2 + W) (cfarOS: lead ++ lagg)
RSP_HD int cfar_kth_largest(const FxArgs &a, const uint32_t *bufM, int start, int W, int k, int start2 = -1) {
  const int bias = 1 << (a.in_w - 1);
  int ans = 0;
  for (int b = a.in_w - 1; b >= 0; --b) {
    const int cand = ans | (1 << b);
    int cnt = 0;
    for (int j = 0; j < W; ++j) cnt += (cfar_x(a, bufM, start + j) + bias >= cand) ? 1 : 0;
    if (start2 >= 0) for (int j = 0; j < W; ++j) cnt += (cfar_x(a, bufM, start2 + j) + bias >= cand) ? 1 : 0;
    if (cnt >= k) ans = cand;
  }
  return ans - bias;
}

// phase D: threshold, peak and output beat of one bin
template <int S>
RSP_HD void cfar_bin(const FxArgs &a, uint32_t row, int n, const uint32_t *bufM, const uint32_t *bufP, uint32_t *gout) {
  constexpr int N = Geo<S>::N;
  const int W = a.W, G = a.G;
  const bool lagg_full = (n + G + W <= N - 1), lead_full = (n - G - W >= 0);
  long long Lv = 0, Rv = 0;
  const bool sums = (a.core == CORE_MEM) || (a.core == CORE_LIS && a.cfar_algorithm == 0);
  if (sums) {
    int sl = 0, sr = 0;
    if (lagg_full) sl = (int)(bufP[pad(n + G + W + 1)] - bufP[pad(n + G + 1)]);
    if (lead_full) sr = (int)(bufP[pad(n - G)] - bufP[pad(n - G - W)]);
    Lv = wrapw(sl, a.sumw) >> a.div_sum;
    Rv = wrapw(sr, a.sumw) >> a.div_sum;
  } else if (a.core == CORE_OS) {
    // cfarOS (CFARUtils.scala:165-171): (lead ++ lagg).sorted((1.5 * referenceCells).toInt), i.e. ascending index floor(1.5 W) of the
    // 2 W reference cells = the (2 W - floor(1.5 W))-th largest; only where both windows exist (the golden pads the edges with 0)
    if (lagg_full && lead_full) Lv = Rv = cfar_kth_largest(a, bufM, n - G - W, W, 2 * W - (3 * W) / 2, n + G + 1);
  } else if (a.core == CORE_LIS) {
    if (lagg_full) Lv = cfar_kth_largest(a, bufM, n + G + 1, W, a.idx_lagg);
    if (lead_full) Rv = cfar_kth_largest(a, bufM, n - G - W, W, a.idx_lead);
  } else { // ASR core: the newest sub_cells of each window (CFARCoreWithASR.scala:285-288)
    const int sub = a.sub_cells;
    int sl = 0, sr = 0;
    if (lagg_full) sl = (int)(bufP[pad(n + G + W + 1)] - bufP[pad(n + G + W - sub + 1)]);
    if (lead_full) sr = (int)(bufP[pad(n - G)] - bufP[pad(n - G - sub)]);
    Lv = wrapw(sl, a.sumw) >> a.div_sum;
    Rv = wrapw(sr, a.sumw) >> a.div_sum;
  }
  const long long go = Lv > Rv ? Lv : Rv, so = Lv < Rv ? Lv : Rv;
  long long by_mode = a.cfar_mode == 0 ? ((Rv + Lv) >> 1) : (a.cfar_mode == 1 ? go : so);
  if (a.core == CORE_OS) by_mode = Lv;   // one statistic: cfarMode does not apply
  if (a.core == CORE_ASR && a.cfar_mode == 3 && lagg_full && lead_full) {
    // CASH: min over sub-windows of max(laggSum_i, leadSum_i), >> log2(subCells) (CFARCoreWithASR.scala:193-209)
    const int sub = a.sub_cells, ns = W / sub;
    long long mn = 0;
    for (int i = 0; i < ns; ++i) {
      const int hl = n + G + W - i * sub, hr = n - G - 1 - i * sub;   // newest cell of sub-window i on each side
      const int sl = wrapw((int)(bufP[pad(hl + 1)] - bufP[pad(hl - sub + 1)]), a.sumw);
      const int sr = wrapw((int)(bufP[pad(hr + 1)] - bufP[pad(hr - sub + 1)]), a.sumw);
      const long long mx = sl > sr ? sl : sr;
      if (i == 0 || mx < mn) mn = mx;
    }
    by_mode = mn >> a.sub_log2;
  }
  long long thr0;
  if (a.core == CORE_MEM) { // one-sided windows at the edges (CFARCoreWithMem.scala:154-171)
    if (!lead_full && !lagg_full) thr0 = 0;
    else if (lagg_full && !lead_full) thr0 = Lv;
    else if (!lagg_full) thr0 = Rv;
    else thr0 = by_mode;
  } else thr0 = (lagg_full && lead_full) ? by_mode : 0;   // CFARCoreWithLis.scala:172-173, CFARCoreWithASR.scala:311-315
  long long thr; int tbp;
  if (a.log_or_linear) {
    const int prod_bp = a.in_bp + a.sc_bp, ctx_bp = (a.in_bp > a.sc_bp ? a.in_bp : a.sc_bp) + 1;
    const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
    thr = trim_rt<long long>(thr0 * (long long)a.scaler, s1, a.cfar_trim);
    tbp = prod_bp - s1;
  } else {
    tbp = a.in_bp > a.sc_bp ? a.in_bp : a.sc_bp;
    thr = (thr0 << (tbp - a.in_bp)) + ((long long)a.scaler << (tbp - a.sc_bp));
  }
  const int cut = cfar_x(a, bufM, n);
  bool pk = ((long long)cut << (tbp - a.in_bp)) > thr;
  if (a.core == CORE_OS && !(lagg_full && lead_full)) { pk = false; thr = 0; }   // the golden's loop runs over windowCells until totalCells - windowCells only; the rest is padded with 0.0 (log mode too)
  if (a.peak_grouping) {
    // neighbours come from the guard registers (CFARCoreWithMem.scala:182-184; CFARCoreWithLis.scala:191-192)
    int left = 0, right = 0;
    if (G != 0) {
      const bool gfull = (G == a.guard_window_size);
      left = (gfull && n + 1 < N) ? cfar_x(a, bufM, n + 1) : 0;
      if (a.core == CORE_MEM) right = (n + G < N) ? cfar_x(a, bufM, n + G) : 0;
      else right = (n - 1 >= 0) ? cfar_x(a, bufM, n - 1) : 0;
    }
    pk = pk && (cut > left) && (cut > right);
  }
  long long t = thr;
  if (tbp > a.thr_bp) t >>= (tbp - a.thr_bp); else if (tbp < a.thr_bp) t <<= (a.thr_bp - tbp);
  const int thr_out = (int)wrapw64(t, a.thr_w);
  if (gout) {
    unsigned long long wd;
    if (a.send_cut) wd = (unsigned long long)(uint32_t)cut & ((1ULL << a.in_w) - 1);
    else wd = (((unsigned long long)(uint32_t)thr_out & ((1ULL << a.thr_w) - 1)) << (a.binw + 1)) |
              (((unsigned long long)n & ((1ULL << a.binw) - 1)) << 1) | (pk ? 1ULL : 0ULL);
    gout[(size_t)row * N + n] = (uint32_t)wd;
  }
  if (pk && a.det_list) {
#if defined(__CUDA_ARCH__)
    const uint32_t slot = atomicAdd(a.det_count, 1u);
#else
    const uint32_t slot = (*a.det_count)++;
#endif
    if (slot < a.det_cap) { Detection d; d.chirp = a.row_base + row; d.bin = (uint32_t)n; d.threshold = thr_out; d.cut = cut; a.det_list[slot] = d; }
  }
}

}  // namespace rsp
