// rsp_device_f32.cuh -- per-thread device code of the fp32 flavour of the chain (the generator elaborated with DspReal,
// generators/sdf-fft/src/main/scala/SDFFFTParams.scala:151-206, run in single precision).  Same phase structure as
// rsp_device.cuh; the result is FFT(x)/N in natural order, JPL / |z|^2 / log2 magnitude in float, and the CFAR cores'
// edge policies with float sums.  Held to 1e-4 relative error against the fp64 oracle, with identical detections.
#pragma once
#include <math.h>
#include <stdint.h>

#include "rsp_device.cuh"

namespace rsp {

struct F2 { float x, y; };

struct F32Args {
  const F2 *in;          // chain / fft input: (re, im)
  const float *in_mag;   // CFAR-only input: one float per bin
  float *out;            // thresholds (CFAR), magnitudes (MAG)
  F2 *out_c;             // FFT-only output
  uint32_t *peak_mask;   // 1 bit per bin, ceil(N/32) words per chirp (nullable)
  uint32_t n_rows, row_base;
  int first_stage, last_stage;
  int bitrev, fft_dir;
  int load_rev;          // DITDecimType without useBitReverse: the beats arrive in bit-reversed order (SDFFFT.scala:254-270); the DspReal
                         // DIT chain computes the same FFT(x)/N as the DIF one, so the DIF dataflow serves both
  const F2 *tw;
  int tw_off[16];
  const float *win;      // window coefficients, read when win_on
  int win_on;
  int sel;
  // CFAR
  int core, guard_window_size, binw, send_cut;
  float scaler, inv_div;
  int log_or_linear, div_sum, peak_grouping, cfar_algorithm, cfar_mode, W, G, idx_lagg, idx_lead, sub_cells, sub_log2;
  Detection *det_list;
  uint32_t *det_count;
  uint32_t det_cap;
};

RSP_HD int padf(int p) { return p + (p >> 4); }   // complex (8-byte) buffers: one pad element per 16

template <int S> struct GeoF {
  static constexpr int N = 1 << S;
  static constexpr int PTS = Geo<S>::PTS, TPC = Geo<S>::TPC, CPC = Geo<S>::CPC, THREADS = Geo<S>::THREADS;
  static constexpr int BUFA = N + (N >> 4) + 2;     // F2 elements
  static constexpr int BUFM = N + (N >> 5) + 2;     // floats
  static constexpr int MASKW = (N + 31) / 32;       // peak-mask words per chirp
};

RSP_HD float jpl_f32(float re, float im) {
  const float a = fabsf(re), b = fabsf(im);
  const float u = a > b ? a : b, v = a > b ? b : a;
  const float x = u + v * 0.125f, y = 0.875f * u + 0.5f * v;
  return x > y ? x : y;
}
RSP_HD float mag_f32(int sel, float re, float im) {
  if (sel == 0) return im;                         // bypass: the CFAR block takes the low half of the beat
  if (sel == 1) return re * re + im * im;
  const float j = jpl_f32(re, im);
  return sel == 2 ? j : log2f(j);
}

template <int S, int K>
RSP_HD void fft_pass_f32(const F32Args &a, int tl, const F2 *gin, F2 *bufA, float *bufM, F2 *gout_c) {
  constexpr int m = fx_pass_size(S, K), i0 = fx_pass_start(S, K), LDl = S - i0 - m, Dl = 1 << LDl;
  constexpr int PTS = Geo<S>::PTS, G = PTS >> m, R = 1 << m, N = 1 << S;
  constexpr bool first = (K == 0), last = (K == fx_n_passes(S) - 1);
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int vt = tl * G + g;
    const int low = vt & (Dl - 1), high = vt >> LDl;
    const int base = (high << (m + LDl)) | low;
    float re[R], im[R];
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int p = base + rr * Dl;
      F2 v;
      if (first) {
        v = gin[a.load_rev ? (int)bitrev_n((uint32_t)p, S) : p];
        if (!a.fft_dir) { const float t = v.x; v.x = v.y; v.y = t; }
        else if (a.win_on) { const float w = a.win[p]; v.x *= w; v.y *= w; }   // SDFFFT.scala:233-234
      }
      else v = bufA[padf(p)];
      re[rr] = v.x; im[rr] = v.y;
    }
#pragma unroll
    for (int s = 0; s < m; ++s) {
      const int i = i0 + s;
      const int dist = 1 << (m - 1 - s);
#pragma unroll
      for (int rr = 0; rr < R; ++rr)
        if (!(rr & dist)) {
          const float sr = re[rr] + re[rr + dist], si = im[rr] + im[rr + dist];
          const float dr = re[rr] - re[rr + dist], di = im[rr] - im[rr + dist];
          re[rr] = sr; im[rr] = si; re[rr + dist] = dr; im[rr + dist] = di;
        }
      if (i == S - 1) continue;
      if ((i & 1) == 0) {
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
          if ((rr & dist) && (rr & (dist >> 1))) { const float r = re[rr]; re[rr] = im[rr]; im[rr] = -r; }
      } else {
        const int L = S - i + 1;
        const int q = 1 << (L - 2);
        const F2 *tab = a.tw + a.tw_off[L];
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
          const int quarter = ((rr >> (m - 1 - s)) & 1) + 2 * ((rr >> (m - s)) & 1);
          if (quarter) {
            const int k = (rr & (dist - 1)) * Dl + low;
            const int slot = quarter == 1 ? 0 : (quarter == 2 ? 1 : 2);
            const F2 w = tab[slot * q + k];
            const float r = re[rr] * w.x - im[rr] * w.y, ii = re[rr] * w.y + im[rr] * w.x;
            re[rr] = r; im[rr] = ii;
          }
        }
      }
    }
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int p = base + rr * Dl;
      if (!last) { F2 v; v.x = re[rr]; v.y = im[rr]; bufA[padf(p)] = v; continue; }
      const int bin = a.bitrev ? (int)bitrev_n((uint32_t)p, S) : p;
      const float sc = 1.0f / (float)N;              // every stage halves (scale-every-stage flavour): exact power of two
      float orr = re[rr] * sc, oi = im[rr] * sc;
      if (!a.fft_dir) { const float t = orr; orr = oi; oi = t; }
      if (a.last_stage == STAGE_FFT) { F2 v; v.x = orr; v.y = oi; gout_c[bin] = v; }   // block-only path: scattered store
      else bufM[pad(bin)] = mag_f32(a.sel, orr, oi);
    }
  }
}

// order-preserving map float -> uint32 (for the bitwise rank search)
RSP_HD uint32_t f32_key(float f) {
  uint32_t b;
#if defined(__CUDA_ARCH__)
  b = __float_as_uint(f);
#else
  memcpy(&b, &f, 4);
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
RSP_HD float f32_unkey(uint32_t k) {
  const uint32_t b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
  float f;
#if defined(__CUDA_ARCH__)
  f = __uint_as_float(b);
#else
  memcpy(&f, &b, 4);
#endif
  return f;
}
RSP_HD int32_t f32_bits(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_int(f);
#else
  int32_t b; memcpy(&b, &f, 4); return b;
#endif
}

RSP_HD float cfar_kth_largest_f32(const float *bufM, int start, int W, int k, int start2 = -1) {   // start2 >= 0: union of two windows (cfarOS)
  uint32_t ans = 0;
  for (int b = 31; b >= 0; --b) {
    const uint32_t cand = ans | (1u << b);
    int cnt = 0;
    for (int j = 0; j < W; ++j) cnt += (f32_key(bufM[pad(start + j)]) >= cand) ? 1 : 0;
    if (start2 >= 0) for (int j = 0; j < W; ++j) cnt += (f32_key(bufM[pad(start2 + j)]) >= cand) ? 1 : 0;
    if (cnt >= k) ans = cand;
  }
  return f32_unkey(ans);
}
RSP_HD float cfar_sum_f32(const float *bufM, int start, int n) {
  float s = 0.f;
  for (int j = 0; j < n; ++j) s += bufM[pad(start + j)];
  return s;
}

// CFAR of the PTS consecutive bins of one thread: sliding window sums kept in registers.  Thresholds go to bufT (padded
// like bufM) and the thread's peak bits are returned, to be written out coalesced by the next phase.
template <int S>
RSP_HD uint32_t cfar_thread_f32(const F32Args &a, uint32_t row, int tl, const float *bufM, float *bufT) {
  constexpr int N = 1 << S, PTS = Geo<S>::PTS;
  const int W = a.W, G = a.G;
  const bool sums = (a.core == CORE_MEM) || (a.core == CORE_LIS && a.cfar_algorithm == 0);
  const int wl = a.core == CORE_ASR ? a.sub_cells : W;   // cells in the running sums
  float sl = 0.f, sr = 0.f;
  uint32_t bits = 0;
  for (int j = 0; j < PTS; ++j) {
    const int n = tl * PTS + j;
    const bool lagg_full = (n + G + W <= N - 1), lead_full = (n - G - W >= 0);
    float Lv = 0.f, Rv = 0.f;
    if (sums || a.core == CORE_ASR) {
      // lagging sum over [n+G+W-wl+1, n+G+W], leading over [n-G-wl, n-G-1], each summed afresh: a sliding float sum (add the
      // entering cell, subtract the leaving one) keeps the rounding error of the largest cell that ever passed through the window,
      // which buries the noise floor behind a strong target (power cells span 7 decades: 1.7e-2 relative error measured on B200)
      if (lagg_full) sl = cfar_sum_f32(bufM, n + G + W - wl + 1, wl);
      if (lead_full) sr = cfar_sum_f32(bufM, n - G - wl, wl);
      Lv = lagg_full ? sl * a.inv_div : 0.f;
      Rv = lead_full ? sr * a.inv_div : 0.f;
    } else if (a.core == CORE_OS) {   // cfarOS (CFARUtils.scala:165-171), see cfar_bin
      if (lagg_full && lead_full) Lv = Rv = cfar_kth_largest_f32(bufM, n - G - W, W, 2 * W - (3 * W) / 2, n + G + 1);
    } else {
      if (lagg_full) Lv = cfar_kth_largest_f32(bufM, n + G + 1, W, a.idx_lagg);
      if (lead_full) Rv = cfar_kth_largest_f32(bufM, n - G - W, W, a.idx_lead);
    }
    const float go = Lv > Rv ? Lv : Rv, so = Lv < Rv ? Lv : Rv;
    float by_mode = a.cfar_mode == 0 ? 0.5f * (Rv + Lv) : (a.cfar_mode == 1 ? go : so);
    if (a.core == CORE_OS) by_mode = Lv;
    if (a.core == CORE_ASR && a.cfar_mode == 3 && lagg_full && lead_full) {
      const int sub = a.sub_cells, ns = W / sub;
      float mn = 0.f;
      for (int i = 0; i < ns; ++i) {
        const float xl = cfar_sum_f32(bufM, n + G + W - i * sub - sub + 1, sub);
        const float xr = cfar_sum_f32(bufM, n - G - 1 - i * sub - sub + 1, sub);
        const float mx = xl > xr ? xl : xr;
        if (i == 0 || mx < mn) mn = mx;
      }
      by_mode = mn / (float)sub;
    }
    float t0;
    if (a.core == CORE_MEM) {
      if (!lead_full && !lagg_full) t0 = 0.f;
      else if (lagg_full && !lead_full) t0 = Lv;
      else if (!lagg_full) t0 = Rv;
      else t0 = by_mode;
    } else t0 = (lagg_full && lead_full) ? by_mode : 0.f;
    float thr = a.log_or_linear ? t0 * a.scaler : t0 + a.scaler;
    const float cut = bufM[pad(n)];
    bool pk = cut > thr;
    if (a.core == CORE_OS && !(lagg_full && lead_full)) { pk = false; thr = 0.f; }
    if (a.peak_grouping) {
      float left = 0.f, right = 0.f;
      if (G != 0) {
        const bool gfull = (G == a.guard_window_size);
        left = (gfull && n + 1 < N) ? bufM[pad(n + 1)] : 0.f;
        if (a.core == CORE_MEM) right = (n + G < N) ? bufM[pad(n + G)] : 0.f;
        else right = (n - 1 >= 0) ? bufM[pad(n - 1)] : 0.f;
      }
      pk = pk && (cut > left) && (cut > right);
    }
    bufT[pad(n)] = thr;
    if (pk) {
      bits |= 1u << j;
      if (a.det_list) {
#if defined(__CUDA_ARCH__)
        const uint32_t slot = atomicAdd(a.det_count, 1u);
#else
        const uint32_t slot = (*a.det_count)++;
#endif
        if (slot < a.det_cap) { Detection d; d.chirp = a.row_base + row; d.bin = (uint32_t)n; d.threshold = f32_bits(thr); d.cut = f32_bits(cut); a.det_list[slot] = d; }
      }
    }
  }
  return bits;
}

}  // namespace rsp
