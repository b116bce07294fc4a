// rsp_fast_f32.cu -- the specialised fused chain kernel of the fp32 flavour (the generator elaborated with DspReal,
// generators/sdf-fft/src/main/scala/SDFFFTParams.scala:151-206, run in single precision): forward radix-2^2 DIF transform
// (SDFChainRadix22.scala:224-352) with natural-order output scaled by 1/N, JPL magnitude (LogMagMux.scala:92-110),
// CFARCoreWithMem in CA / GO / SO mode with a linear or log threshold (CFARCoreWithMem.scala:102-189); threshold per bin +
// 1 peak bit per bin out.  Anything else runs on chain_f32_kernel (rsp_kernels.cuh); both are held to the same tolerance
// against the fp64 oracle (tests/test_gpu_f32_gen.py).
//
// This flavour is the one that can reach the HBM roofline (12.125 B and ~55 instructions per sample), so the kernel is
// written for issue slots and shared-memory wavefronts:
//  * a point (re, im) lives in one 64-bit register pair and the butterflies are Blackwell's packed fp32 instructions
//    (FADD2 / FMUL2 / FFMA2, PTX add/mul/fma.f32x2): one instruction per complex add, two per complex twiddle product
//    ((c, s) * re then (-s, c) * im + acc, the table holds (c, s, -s, c))
//  * the x(-j) rotation is never executed: the butterfly that consumes a rotated point is written out on the swapped lanes
//  * three (four) register passes of up to 4 radix-2 stages; between passes the points cross shared memory once, at
//    index p + 2 (p >> 4) + 4 (p >> 6), which makes the 64-bit and 128-bit access patterns of the passes conflict-free
//  * the bit reversal is the scatter of the magnitudes; the CFAR window sums are differences of one prefix-sum array that
//    reuses the exchange buffer; bins are taken interleaved so that loads have immediate offsets and stores are coalesced
//  * the peak bits of 32 consecutive bins are one warp ballot.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "rsp_fast.cuh"

namespace rsp {

typedef unsigned long long u64;

struct FastF32Args {
  const u64 *in;           // (re, im) float pairs
  float *out;              // thresholds (nullable)
  uint32_t *peak_mask;     // N/32 words per row (nullable)
  uint32_t n_rows, row_base;
  const float *win;        // window coefficients or null (FFTParams.windowFunc, SDFFFT.scala:53-66,233-234)
  const float4 *tw;        // (c, s, -s, c), c + j s = exp(-2 pi j t / stageN); per stageN three runs [W^2k | W^k | W^3k]
  int tw_off[16];
  int W, G;
  float k_both, k_one;     // linear: threshold = comb(L, R) * k_both, one-sided L * k_one; log: comb * k_both + scaler
  float scaler, inv_n;
  Detection *det_list;
  uint32_t *det_count;
  uint32_t det_cap;
  int pg_full;             // peak grouping: -1 off; else 1 = guardCells equals guardWindowSize (the left neighbour is read), 0 = it reads as 0
  int l2_prefetch;         // 1: one thread asks the L2 for the CTA's next row while this one is transformed (cp.async.bulk.prefetch.L2)
};

// Points per thread: 16, or two sets of 16 (NQ = 2) from N = 2^RSP_F32_NQ2_SMIN on.  N = 16384 at 16 points per thread is one CTA of 1024
// threads, which leaves ptxas 64 registers and 116 bytes of spills; at 32 points it is 512 threads with 128 registers.  A pass still
// works on 16 points at a time, the thread just owns twice as many groups (and two blocks of 16 bins in the CFAR).
#ifndef RSP_F32_REGPF
#define RSP_F32_REGPF 1   // next row's beats into registers behind the first pass (A/B knob)
#endif
#ifndef RSP_F32_NQ2_SMIN
#define RSP_F32_NQ2_SMIN 14
#endif
template <int S> struct FGeoF {
  static constexpr int N = 1 << S;
  static constexpr int NQ = S >= RSP_F32_NQ2_SMIN ? 2 : 1;
  static constexpr int LP = NQ == 2 ? 5 : 4;
  static constexpr int PTS = 16 * NQ;
  static constexpr int TPC = N / PTS;                         // threads per chirp
  static constexpr int CPC = TPC >= 128 ? 1 : 128 / TPC;      // chirps per CTA
  static constexpr int THREADS = TPC * CPC;
  static constexpr int XB = N + (N >> 3) + (N >> 4) + 8;      // u64 elements of the exchange buffer, index xidx(p)
  static constexpr int MG = N + (N >> 3) + 8;                 // floats of the magnitude array, index midx(n)
  static constexpr size_t ROW_BYTES = (size_t)XB * 8 + (size_t)MG * 4;
  static constexpr int NW = TPC / 32;                         // warps per chirp
  static constexpr int MASKW = N / 32;
  static constexpr int JSTEP = TPC + (TPC >> 3);              // midx(n + TPC) - midx(n)
};

// exchange-buffer index of point p: one 16-byte pad per 16 points, and (except at N = 4096, whose last pass reads 16 neighbouring
// points per thread and is conflict-free without it) one 32-byte pad per 64 (tests/test_kernel_layout.py models every pass)
template <int S> __host__ __device__ constexpr int xidx(int p) { return p + ((p >> 4) << 1) + (S == 12 ? 0 : ((p >> 6) << 2)); }
__device__ __forceinline__ int midx(int n) { return n + ((n >> 5) << 2); }

__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(u64 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
// Packed or scalar arithmetic: measured on B200 (scripts/ubench/fp2.cu, profiles/r01_f_ubench_fp2.txt) FADD2/FFMA2 issue at
// 1.1 warp-instructions/clk/SM against 2.7 for FADD/FFMA, i.e. a packed instruction saves an issue slot but costs ~25% more
// FMA-pipe time than the two scalar ones.  RSP_F32_PACKED selects which of the two this kernel is built from.
#ifndef RSP_F32_PACKED
#define RSP_F32_PACKED 1
#endif
#if RSP_F32_PACKED
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
#else
__device__ __forceinline__ u64 add2(u64 a, u64 b) { float a0, a1, b0, b1; upk2(a, a0, a1); upk2(b, b0, b1); return pk2(a0 + b0, a1 + b1); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { float a0, a1, b0, b1; upk2(a, a0, a1); upk2(b, b0, b1); return pk2(a0 - b0, a1 - b1); }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { float a0, a1, b0, b1; upk2(a, a0, a1); upk2(b, b0, b1); return pk2(a0 * b0, a1 * b1); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  float a0, a1, b0, b1, c0, c1; upk2(a, a0, a1); upk2(b, b0, b1); upk2(c, c0, c1);
  return pk2(fmaf(a0, b0, c0), fmaf(a1, b1, c1));
}
#endif
// a <- a + c, c <- a - c
__device__ __forceinline__ void bf_plain(u64 &a, u64 &c) { const u64 s = add2(a, c), d = sub2(a, c); a = s; c = d; }
// the same with c standing for -j c = (c.im, -c.re) (SDFChainRadix22.scala:319-337: the rotation after an even stage)
__device__ __forceinline__ void bf_rot(u64 &a, u64 &c) {
  float ar, ai, cr, ci;
  upk2(a, ar, ai); upk2(c, cr, ci);
  a = pk2(ar + ci, ai - cr);
  c = pk2(ar - ci, ai + cr);
}
// x <- x * (c + j s), w = (c, s, -s, c)
__device__ __forceinline__ u64 cmul(u64 x, const float4 w) {
  float xr, xi;
  upk2(x, xr, xi);
  return fma2(pk2(xi, xi), pk2(w.z, w.w), mul2(pk2(xr, xr), pk2(w.x, w.y)));
}
// x <- x * w with w given as its two operand pairs A = (c, s), B = (-s, c)
__device__ __forceinline__ u64 cmul2(u64 x, u64 A, u64 B) {
  float xr, xi;
  upk2(x, xr, xi);
  return fma2(pk2(xi, xi), B, mul2(pk2(xr, xr), A));
}
// JPL magnitude approximation (LogMagMux.scala:92-110) of two points at once, times `sc` (a power of two: the 1/N of the
// scale-every-stage transform): max(u + v/8, 7u/8 + v/2) with u = max(|re|, |im|), v = min(|re|, |im|)
__device__ __forceinline__ void jpl2_scaled(u64 z0, u64 z1, float sc, float &m0, float &m1) {
  float r0, i0, r1, i1;
  upk2(z0, r0, i0); upk2(z1, r1, i1);
  const u64 u = pk2(fmaxf(fabsf(r0), fabsf(i0)), fmaxf(fabsf(r1), fabsf(i1)));
  const u64 v = pk2(fminf(fabsf(r0), fabsf(i0)), fminf(fabsf(r1), fabsf(i1)));
  const u64 x = fma2(v, pk2(0.125f, 0.125f), u);
  const u64 y = fma2(v, pk2(0.5f, 0.5f), mul2(u, pk2(0.875f, 0.875f)));
  float x0, x1, y0, y1;
  upk2(x, x0, x1); upk2(y, y0, y1);
  const u64 mm = mul2(pk2(fmaxf(x0, y0), fmaxf(x1, y1)), pk2(sc, sc));
  upk2(mm, m0, m1);
}


// One pass = m radix-2 stages in registers on 2^m points Dl apart (the decomposition of rsp_device.cuh, fx_pass_*).
// Thread tl of the chirp works on the groups vt = g * TPC + tl, so a warp touches consecutive points in every pass.
// Every shared-memory address is one run-time base plus a compile-time offset: with gbase(v) = the first point of group v (m zero bits
// inserted at bit LDl of v), gbase(g * TPC + tl) = gbase(tl) + gbase(g * TPC) and xidx(gbase(tl) + gbase(g * TPC) + rr * Dl) =
// xidx(gbase(tl)) + xidx(gbase(g * TPC) + rr * Dl) for all the shapes used here (checked exhaustively, tests/test_kernel_layout.py).
template <int m, int LDl> __host__ __device__ constexpr int gbase(int v) { return ((v >> LDl) << (m + LDl)) | (v & ((1 << LDl) - 1)); }
template <int S, int K>
__device__ __forceinline__ void f32_pass(const FastF32Args &a, int tl, const u64 (&pre)[FGeoF<S>::PTS], u64 *xb, float *mg) {
  using Gm = FGeoF<S>;
  constexpr int m = fx_pass_size(S, K), i0 = fx_pass_start(S, K), LDl = S - i0 - m, Dl = 1 << LDl;
  constexpr int PTS = Gm::PTS, G = PTS >> m, R = 1 << m, TPC = Gm::TPC, LP = Gm::LP;
  constexpr bool first = (K == 0), last = (K == fx_n_passes(S) - 1);
  static_assert((i0 & 1) == 0, "passes start on even stages");
  static_assert(!last || Dl == 1, "the last pass works on neighbouring points");
  float mags[last ? PTS : 1];
  const int low = tl & (Dl - 1);
  const int base0 = ((tl >> LDl) << (m + LDl)) | low;        // gbase(tl); group g starts at base0 + gbase(g * TPC)
  u64 *xp = xb + xidx<S>(base0);
#pragma unroll
  for (int g = 0; g < G; ++g) {
    u64 z[R];
    if (first) {
#pragma unroll
      for (int rr = 0; rr < R; ++rr) z[rr] = pre[g + G * rr];   // m = 4, G = NQ groups, p = tl + g * TPC + rr * (NQ * TPC)
    } else if (last) {                                        // 2^m neighbouring points: 128-bit loads
#pragma unroll
      for (int rr = 0; rr < R; rr += 2) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(&xp[xidx<S>(gbase<m, LDl>(g * TPC) + rr)]);
        z[rr] = v.x; z[rr + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int rr = 0; rr < R; ++rr) z[rr] = xp[xidx<S>(gbase<m, LDl>(g * TPC) + rr * Dl)];
    }
#pragma unroll
    for (int s = 0; s < m; ++s) {
      const int i = i0 + s;
      const int dist = 1 << (m - 1 - s);
#pragma unroll
      for (int rr = 0; rr < R; ++rr)
        if (!(rr & dist)) {
          // odd stage: the partner was the difference output of the previous stage's butterfly (bit 2 dist of rr) and sits
          // in the last quarter of its 4 dist block -> it carries the pending x(-j)
          const bool rot = (i & 1) && (rr & (dist << 1));
          if (rot) bf_rot(z[rr], z[rr + dist]); else bf_plain(z[rr], z[rr + dist]);
        }
      if (i == S - 1 || (i & 1) == 0) continue;
      const int L = S - i + 1;
      const int q = 1 << (L - 2);
      // one table read per distinct k: W^k with both operand pairs A = (c, s), B = (-s, c); W^2k = W^k W^k and
      // W^3k = W^2k W^k follow by the same packed product (B' = j W' = (j W) W), which saves two thirds of the loads
      const float4 *tab = a.tw + a.tw_off[L] + q + low + ((g * TPC) & (Dl - 1));   // the W^k run; k = (vt mod Dl) + (rr mod dist) Dl
#pragma unroll
      for (int kk = 0; kk < dist; ++kk) {
        const float4 w1 = __ldg(&tab[kk * Dl]);
        const u64 A1 = pk2(w1.x, w1.y), B1 = pk2(w1.z, w1.w);
        const u64 A2 = cmul2(A1, A1, B1), B2 = cmul2(B1, A1, B1);
        const u64 A3 = cmul2(A2, A1, B1), B3 = cmul2(B2, A1, B1);
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
          if ((rr & (dist - 1)) != kk) continue;
          const int quarter = ((rr >> (m - 1 - s)) & 1) + 2 * ((rr >> (m - s)) & 1);
          if (quarter == 1) z[rr] = cmul2(z[rr], A2, B2);
          else if (quarter == 2) z[rr] = cmul2(z[rr], A1, B1);
          else if (quarter == 3) z[rr] = cmul2(z[rr], A3, B3);
        }
      }
    }
    if (!last) {
#pragma unroll
      for (int rr = 0; rr < R; ++rr) xp[xidx<S>(gbase<m, LDl>(g * TPC) + rr * Dl)] = z[rr];
    } else {
#pragma unroll
      for (int rr = 0; rr < R; rr += 2) jpl2_scaled(z[rr], z[rr + 1], a.inv_n, mags[g * R + rr], mags[g * R + rr + 1]);
    }
  }
  if (last) {
    // bin of point p = (vt << m) + rr is bitrev_S(p) = rev_m(rr) << (S - m) | rev_{S-LP}(tl) << (LP - m) | rev_{LP-m}(g):
    // for a fixed rr the G groups of the thread are G neighbouring bins -> one G-float store per rr
    const uint32_t rtl = (__brev((uint32_t)tl) >> (32 - (S - LP))) << (LP - m);
    float *mp = mg + midx((int)rtl);
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      constexpr int SH = S - m;
      const int rrr = (int)((__brev((uint32_t)rr) >> (32 - m)) << SH);   // compile-time, a multiple of 32
      float *dst = mp + rrr + ((rrr >> 5) << 2);
      if constexpr (G >= 4) {
        float o[G];
#pragma unroll
        for (int g = 0; g < G; ++g) o[(int)(__brev((uint32_t)g) >> (32 - (LP - m)))] = mags[g * R + rr];
#pragma unroll
        for (int q = 0; q < G; q += 4) *reinterpret_cast<float4 *>(dst + q) = make_float4(o[q], o[q + 1], o[q + 2], o[q + 3]);
      } else if constexpr (G == 2) *reinterpret_cast<float2 *>(dst) = make_float2(mags[0 * R + rr], mags[1 * R + rr]);
      else *dst = mags[rr];
    }
  }
}

template <int S, int K>
__device__ __forceinline__ void f32_passes(const FastF32Args &a, int tl, const u64 (&pre)[FGeoF<S>::PTS], u64 *xb, float *mg) {
  static_assert(fx_pass_size(S, 0) == 4, "the first pass takes the prefetched points of the thread, 16 at a time");
  f32_pass<S, K>(a, tl, pre, xb, mg);
  __syncthreads();
  if constexpr (K + 1 < fx_n_passes(S)) f32_passes<S, K + 1>(a, tl, pre, xb, mg);
}

// ---------------------------------------------------------------------------------------------------------------------
// CFAR (CFARCoreWithMem.scala:102-189): lagging window of bin n = [n+G+1, n+G+W], leading window = [n-G-W, n-G-1].

// threshold from the two window sums, explicit roundings so that the detection path reproduces the hot loop bit for bit
template <int CMODE, bool LINEAR, typename A>
__device__ __forceinline__ float f32_threshold(const A &a, float L, float R, bool lagg_full, bool lead_full) {
  float comb = CMODE == 0 ? __fadd_rn(L, R) : (CMODE == 1 ? fmaxf(L, R) : fminf(L, R));
  float k = a.k_both;
  if (!(lagg_full && lead_full)) { comb = lead_full ? R : (lagg_full ? L : 0.f); k = a.k_one; }
  return LINEAR ? __fmul_rn(comb, k) : __fmaf_rn(comb, k, a.scaler);
}

// detections of one thread: bit j of `mask` = peak of bin tl + j * TPC.  Rare path, recomputes from shared memory.
struct FastF32DetArgs { int W, G; float k_both, k_one, scaler; uint32_t chirp, mask, det_cap; Detection *det_list; uint32_t *det_count; };
template <int S, int CMODE, bool LINEAR>
__device__ __noinline__ void f32_detections(const FastF32DetArgs a, int tl, const float *QP, const float *AS, const float *X) {
  constexpr int N = 1 << S, TPC = FGeoF<S>::TPC;
  for (int j = 0; j < FGeoF<S>::PTS; ++j) {
    if (!((a.mask >> j) & 1u)) continue;
    const int n = tl + j * TPC, W = a.W, G = a.G;
    const bool lagg_full = (n + G + W <= N - 1), lead_full = (n - G - W >= 0);
    const int hi = min(n + G + W + 1, N), lo = max(n - G - W, 0);
    const float L = __fadd_rn(AS[hi - W], QP[hi]), R = __fadd_rn(AS[lo], QP[lo + W]);
    const float thr = f32_threshold<CMODE, LINEAR>(a, L, R, lagg_full, lead_full);
    const uint32_t slot = atomicAdd(a.det_count, 1u);
    if (slot < a.det_cap) {
      Detection d; d.chirp = a.chirp; d.bin = (uint32_t)n; d.threshold = __float_as_int(thr); d.cut = __float_as_int(X[midx(n)]);
      a.det_list[slot] = d;
    }
  }
}

// Resident CTAs per SM the register budget is cut for.  Shared memory already limits N = 4096 to three CTAs, N = 8192 and 16384 to one;
// below that the cap is a trade of warps for registers: at 7 CTAs of 128 threads ptxas has 72 registers and spills the 16-point
// butterflies, at 5 it has 96 and does not -- measured on B200 (profiles/r02_q_f32_variants.txt): C3 x 16 239 -> 275 Gsamples/s
// from 7 -> 5, N = 4096 195 -> 211 from 4 -> 3.
#ifndef RSP_F32_MAXB
#define RSP_F32_MAXB 5
#endif
#ifndef RSP_F32_MAXB12
#define RSP_F32_MAXB12 2
#endif
#ifndef RSP_F32_MAXB13
#define RSP_F32_MAXB13 1
#endif
template <int S> constexpr int f32_min_blocks() {
  constexpr int by_threads = 1024 / FGeoF<S>::THREADS;
  constexpr int by_smem = (int)((size_t)228 * 1024 / (FGeoF<S>::ROW_BYTES * FGeoF<S>::CPC + 1024));   // 1 KiB per CTA is the system's
  constexpr int m = by_threads < by_smem ? by_threads : by_smem;
  constexpr int cap = FGeoF<S>::NQ == 2 ? 8 : (S == 13 ? RSP_F32_MAXB13 : (S == 12 ? RSP_F32_MAXB12 : RSP_F32_MAXB));
  return m > cap ? cap : (m < 1 ? 1 : m);
}

// EJ = number of leading / trailing iterations j whose bins may miss a window (ceil((W + G) / TPC), checked by the host)
// PG: the peak-grouping instantiation (a run-time register; its own kernels so that the common ones keep their register budget)
template <int S, int CMODE, bool LINEAR, int EJ, bool PG = false>
__global__ void __launch_bounds__(FGeoF<S>::THREADS, f32_min_blocks<S>())
chain_f32_fast_kernel(const FastF32Args a) {
  using Gm = FGeoF<S>;
  constexpr int N = Gm::N, TPC = Gm::TPC, CPC = Gm::CPC, PTS = Gm::PTS, JSTEP = Gm::JSTEP, NQ = Gm::NQ;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  u64 *xb = reinterpret_cast<u64 *>(smem_raw + (size_t)c * Gm::ROW_BYTES);
  float *mg = reinterpret_cast<float *>(xb + Gm::XB);
  // QP, AS and the block totals (+ 8 zeros for the blocks past the end of the row) reuse the exchange buffer, dead after the last FFT pass
  float *QP = reinterpret_cast<float *>(xb), *AS = QP + N + 16, *tsum = AS + N;
  static_assert((size_t)(2 * N + 16 + NQ * TPC + 8) * 4 <= (size_t)Gm::XB * 8, "the CFAR arrays fit the exchange buffer");
  // the leading borders of the first iterations are negative: only dereferenced from j >= EJ on, where the index is >= 0
  const float *pA = AS + (tl + a.G + 1), *pB = QP + (tl + a.G + a.W + 1), *pC = AS + (tl - a.G - a.W), *pD = QP + (tl - a.G);
  const int nblk = a.W >> 4;
  const float *pX = mg + midx(tl);
  // the thread's blocks of 16 neighbouring bins: block tl + h TPC, h < NQ; midx(16 b + k) = midx(16 b) + k
  const float *pX16 = mg + midx(tl * 16);
  float *pP16 = QP + tl * 16, *pS16 = AS + tl * 16;
  constexpr int XH = 16 * TPC + ((16 * TPC) >> 3);   // midx(16 (tl + TPC)) - midx(16 tl)
  const int prot = (tl >> 1) & 3;
  const uint32_t stride = gridDim.x * CPC;
  // The thread's points of a row, p = tl + rr * TPC (warp-contiguous 256-byte runs), are only used by the first pass: the next row's are
  // loaded into the same registers right behind it and land while the other passes and the CFAR of this row run (from N = 4096 on, where
  // few warps are resident: +1 % there, -2 % below; profiles/r02_q_f32_variants.txt).  The L2 prefetch runs one row further ahead.
  constexpr bool PF = RSP_F32_REGPF && S >= 12;
  u64 pre[PTS];
  auto load_row = [&](uint32_t r0, int g_lo, int g_hi) {   // the points of the first-pass groups g_lo .. g_hi - 1 (pre[g + NQ * rr])
    const uint32_t r = r0 + c < a.n_rows ? r0 + c : a.n_rows - 1;
    const u64 *gin = a.in + (size_t)r * N + tl;
#pragma unroll
    for (int q = 0; q < PTS; ++q) if (q % NQ >= g_lo && q % NQ < g_hi) pre[q] = __ldg(&gin[q * TPC]);
  };
  auto l2_row = [&](uint32_t r0) {
    if (a.l2_prefetch && tl == 0 && r0 + c < a.n_rows) {
      const u64 *nxt = a.in + (size_t)(r0 + c) * N;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(nxt), "r"(8 * N) : "memory");
    }
  };
  // with 32 points per thread only the first group's 16 are taken ahead (64 more registers would spill); the others follow at the top of the row
  if (PF && blockIdx.x * CPC < a.n_rows) { load_row(blockIdx.x * CPC, 0, 1); l2_row(blockIdx.x * CPC + stride); }
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += stride) {
    const bool active = row0 + c < a.n_rows;
    const uint32_t row = active ? row0 + c : a.n_rows - 1;
    load_row(row0, PF ? 1 : 0, NQ);
    l2_row(row0 + (PF ? 2 : 1) * stride);
    if (a.win != nullptr) {
      const float *gw = a.win + tl;
#pragma unroll
      for (int rr = 0; rr < PTS; ++rr) { const float w = __ldg(&gw[rr * TPC]); pre[rr] = mul2(pre[rr], pk2(w, w)); }
    }
    f32_pass<S, 0>(a, tl, pre, xb, mg);
    if (PF && row0 + stride < a.n_rows) load_row(row0 + stride, 0, 1);
    __syncthreads();
    f32_passes<S, 1>(a, tl, pre, xb, mg);
    // window sums without a row-wide prefix (which would lose the noise floor behind a strong target in fp32): bins are
    // cut into the threads' blocks of 16; QP[n] = sum of the block of n before n, AS[n] = sum of the block of n from n on
    // plus the W/16 - 1 blocks that follow.  A window [a, a + W) with W a multiple of 16 is AS[a] + QP[a + W]: additions
    // of non-negative terms only, every partial sum no larger than the window's own.
    float x[NQ][16];
#pragma unroll
    for (int h = 0; h < NQ; ++h) {
#pragma unroll
      for (int k = 0; k < 16; k += 4) {
        const float4 v = *reinterpret_cast<const float4 *>(&pX16[h * XH + k]);
        x[h][k] = v.x; x[h][k + 1] = v.y; x[h][k + 2] = v.z; x[h][k + 3] = v.w;
      }
      float t = x[h][15];
#pragma unroll
      for (int k = 14; k >= 0; --k) t += x[h][k];
      tsum[tl + h * TPC] = t;
    }
    if (tl < 8) tsum[NQ * TPC + tl] = 0.f;
    __syncthreads();
#pragma unroll
    for (int h = 0; h < NQ; ++h) {
      float follow = 0.f;
      for (int i = 1; i < nblk; ++i) follow += tsum[tl + h * TPC + i];
      float sv[16], pv[16];
      sv[15] = x[h][15];
#pragma unroll
      for (int k = 14; k >= 0; --k) sv[k] = sv[k + 1] + x[h][k];
      pv[0] = 0.f;
#pragma unroll
      for (int k = 1; k < 16; ++k) pv[k] = pv[k - 1] + x[h][k - 1];
#pragma unroll
      for (int k = 0; k < 16; ++k) sv[k] += follow;
      // The block's 16 values are 64 bytes: lanes 64 bytes apart would hit two 16-byte bank groups only, so lane l stores
      // its four quads in the order rotated by (l >> 1) & 3, which spreads the 8 lanes of a quarter warp over all 8 groups.
      const bool r1 = (tl >> 1) & 1, r2 = (tl >> 2) & 1;
#pragma unroll
      for (int arr = 0; arr < 2; ++arr) {
        const float *src = arr ? sv : pv;
        float *dst = (arr ? pS16 : pP16) + h * 16 * TPC;
        float4 f4[4], g4[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) f4[k] = make_float4(src[4 * k], src[4 * k + 1], src[4 * k + 2], src[4 * k + 3]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 p = f4[k], n = f4[(k + 1) & 3];
          g4[k] = make_float4(r1 ? n.x : p.x, r1 ? n.y : p.y, r1 ? n.z : p.z, r1 ? n.w : p.w);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 p = g4[k], n = g4[(k + 2) & 3];
          const float4 v = make_float4(r2 ? n.x : p.x, r2 ? n.y : p.y, r2 ? n.z : p.z, r2 ? n.w : p.w);
          *reinterpret_cast<float4 *>(&dst[((k + prot) & 3) * 4]) = v;
        }
      }
    }
    if (tl == 0) QP[N] = 0.f;
    __syncthreads();
    float *gout = a.out ? a.out + (size_t)row * N + tl : nullptr;
    uint32_t *gmask = a.peak_mask ? a.peak_mask + (size_t)row * Gm::MASKW + (tl >> 5) : nullptr;
    uint32_t mask = 0;
    auto bins = [&](auto pgc) {
#pragma unroll
    for (int j = 0; j < PTS; ++j) {
      float L, R;
      bool lagg_full = true, lead_full = true;
      if (j < EJ || j >= PTS - EJ) {
        const int n = tl + j * TPC, W = a.W, G = a.G;
        lagg_full = (n + G + W <= N - 1); lead_full = (n - G - W >= 0);
        const int hi = min(n + G + W + 1, N), lo = max(n - G - W, 0);
        L = __fadd_rn(AS[hi - W], QP[hi]); R = __fadd_rn(AS[lo], QP[lo + W]);
      } else {
        L = __fadd_rn(pA[j * TPC], pB[j * TPC]); R = __fadd_rn(pC[j * TPC], pD[j * TPC]);
      }
      const float thr = f32_threshold<CMODE, LINEAR>(a, L, R, lagg_full, lead_full);
      bool pk = pX[j * JSTEP] > thr;
      if (decltype(pgc)::value) {   // peakGrouping: both neighbours from the lagging guard (CFARCoreWithMem.scala:182-184)
        const int n = tl + j * TPC, n1 = n + 1, ng = n + a.G;
        const float cut = pX[j * JSTEP];
        const float left = (a.pg_full && n1 < N) ? mg[midx(n1)] : 0.f, right = ng < N ? mg[midx(ng)] : 0.f;
        pk = pk && cut > left && cut > right;
      }
      if (gout) gout[j * TPC] = thr;
      const uint32_t word = __ballot_sync(0xffffffffu, pk);
      if (gmask && (tl & 31) == 0) gmask[j * (TPC / 32)] = word;
      mask |= (pk ? 1u : 0u) << j;
    }
    };
    bins(std::integral_constant<bool, PG>{});
    if (mask != 0 && a.det_list != nullptr && active) {
      const FastF32DetArgs da = {a.W, a.G, a.k_both, a.k_one, a.scaler, a.row_base + row, mask, a.det_cap, a.det_list, a.det_count};
      f32_detections<S, CMODE, LINEAR>(da, tl, QP, AS, mg);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// host side
static int f32_edge_iters(int S, int W, int G) { const int tpc = (1 << S) / (S >= RSP_F32_NQ2_SMIN ? 32 : 16); return (W + G + tpc - 1) / tpc; }

bool fast_f32_chain_supported(const Plan &pl, const F32Args &a) {
  if (pl.first_stage != STAGE_FFT || pl.last_stage != STAGE_CFAR) return false;
  if (pl.S < 10 || pl.S > RSP_FAST_SMAX) return false;
  if (!a.bitrev || a.load_rev || a.fft_dir != 1 || a.sel != 2) return false;
  if (a.core != CORE_MEM || a.send_cut) return false;
  if (a.G < 1 || a.W < 1 || a.div_sum < 0 || a.div_sum > 30) return false;
  if ((a.W & 15) || a.W > 128) return false;   // the block-of-16 window sums (see the kernel)
  if (f32_edge_iters(pl.S, a.W, a.G) > (pl.S == 10 ? 2 : 1)) return false;
  return a.out != nullptr || a.peak_mask != nullptr || a.det_list != nullptr;
}

template <int S, int CMODE, bool LINEAR, int EJ, bool PG = false>
static rsp_status launch_f32_fast_pg(const LaunchCtx &c, const FastF32Args &fa, cudaStream_t st, const char **err) {
  using Gm = FGeoF<S>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)Gm::CPC * Gm::ROW_BYTES;
  auto kernel = chain_f32_fast_kernel<S, CMODE, LINEAR, EJ, PG>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, Gm::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (fa.n_rows + Gm::CPC - 1) / Gm::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, Gm::THREADS, smem, st>>>(fa);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_f32_fast_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}

template <int S, int CMODE, bool LINEAR, int EJ>
static rsp_status launch_f32_fast_inst(const LaunchCtx &c, const FastF32Args &fa, cudaStream_t st, const char **err) {
  return fa.pg_full >= 0 ? launch_f32_fast_pg<S, CMODE, LINEAR, EJ, true>(c, fa, st, err) : launch_f32_fast_pg<S, CMODE, LINEAR, EJ, false>(c, fa, st, err);
}

template <int S, int EJ>
static rsp_status launch_f32_fast_modes(int cmode, bool linear, const LaunchCtx &c, const FastF32Args &fa, cudaStream_t st, const char **err) {
  if (linear) switch (cmode) {
    case 0: return launch_f32_fast_inst<S, 0, true, EJ>(c, fa, st, err);
    case 1: return launch_f32_fast_inst<S, 1, true, EJ>(c, fa, st, err);
    default: return launch_f32_fast_inst<S, 2, true, EJ>(c, fa, st, err);
  }
  switch (cmode) {
    case 0: return launch_f32_fast_inst<S, 0, false, EJ>(c, fa, st, err);
    case 1: return launch_f32_fast_inst<S, 1, false, EJ>(c, fa, st, err);
    default: return launch_f32_fast_inst<S, 2, false, EJ>(c, fa, st, err);
  }
}

rsp_status launch_fast_f32_chain(const Plan &pl, const LaunchCtx &ctx, const F32Args &a, const float4 *tw4, cudaStream_t st, const char **err) {
  FastF32Args fa;
  fa.in = reinterpret_cast<const u64 *>(a.in); fa.out = a.out; fa.peak_mask = a.peak_mask; fa.n_rows = a.n_rows; fa.row_base = a.row_base;
  fa.tw = tw4; fa.win = a.win_on ? a.win : nullptr;
  for (int L = 0; L < 16; ++L) fa.tw_off[L] = a.tw_off[L];
  fa.W = a.W; fa.G = a.G; fa.scaler = a.scaler; fa.inv_n = 1.0f / (float)(1 << pl.S);
  const int cmode = a.cfar_mode == 0 ? 0 : (a.cfar_mode == 1 ? 1 : 2);   // MuxLookup default: smallestOf (CFARCoreWithMem.scala:138-141)
  const bool linear = a.log_or_linear != 0;
  const float both = cmode == 0 ? 0.5f * a.inv_div : a.inv_div;            // CA: (L + R) / 2 (CFARCoreWithMem.scala:135-137)
  fa.k_both = linear ? both * a.scaler : both;
  fa.k_one = linear ? a.inv_div * a.scaler : a.inv_div;
  fa.det_list = a.det_list; fa.det_count = a.det_count; fa.det_cap = a.det_cap;
  { static int l2p = -1; if (l2p < 0) { const char *e = std::getenv("RSP_FAST_L2_PREFETCH"); l2p = e ? std::atoi(e) : 1; }
    fa.l2_prefetch = (l2p != 0 && (reinterpret_cast<uintptr_t>(a.in) & 15u) == 0) ? 1 : 0; }
  fa.pg_full = a.peak_grouping ? (a.G == a.guard_window_size ? 1 : 0) : -1;
  const int ej = f32_edge_iters(pl.S, a.W, a.G);
  switch (pl.S) {
  case 10: return ej == 2 ? launch_f32_fast_modes<10, 2>(cmode, linear, ctx, fa, st, err) : launch_f32_fast_modes<10, 1>(cmode, linear, ctx, fa, st, err);
  case 11: return launch_f32_fast_modes<11, 1>(cmode, linear, ctx, fa, st, err);
  case 12: return launch_f32_fast_modes<12, 1>(cmode, linear, ctx, fa, st, err);
  case 13: return launch_f32_fast_modes<13, 1>(cmode, linear, ctx, fa, st, err);
  case 14: return launch_f32_fast_modes<14, 1>(cmode, linear, ctx, fa, st, err);
  default: *err = "no fast fp32 variant for this size"; return RSP_EUNSUPPORTED;
  }
}

}  // namespace rsp
