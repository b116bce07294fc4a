/*
 * rsp_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY; see rsp_oracle.h for the parity status).
 *
 * Plain-C restatement of the reference generators' arithmetic.  Every function cites the reference
 * file:line it follows (paths relative to /root/reference/).  Semantics of the un-vendored dsptools
 * typeclass operations (S1..S11) are listed in DESIGN.md and SURVEY.md Appendix A.0.
 */
#include "rsp_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------------------------------
 * primitives
 * ---------------------------------------------------------------------------------------------- */

/* sign-extend the low `width` bits (chisel `:=` into a narrower SInt/FixedPoint drops MSBs, S6/S11) */
int64_t rspo_wrap(int64_t v, int width) {
  if (width >= 64) return v;
  uint64_t m = (width == 0) ? 0 : ((~0ULL) >> (64 - width));
  uint64_t u = (uint64_t)v & m;
  if (width > 0 && (u >> (width - 1)) & 1) u |= ~m;
  return (int64_t)u;
}

/* v / 2^shift under a dsptools TrimType (trimBinary / div2, S3/S4) */
int64_t rspo_trim(int64_t v, int shift, int mode) {
  if (shift <= 0 || mode == RSPO_TRIM_NO_TRIM) return v;
  const int64_t one = 1;
  const int64_t half = one << (shift - 1);
  const int64_t mask = (one << shift) - 1;
  const int64_t q = v >> shift; /* floor */
  const int64_t r = v & mask;   /* 0 <= r < 2^shift */
  switch (mode) {
  case RSPO_TRIM_ROUND_HALF_UP: return (v + half) >> shift;
  case RSPO_TRIM_FLOOR: return q;
  case RSPO_TRIM_CONVERGENT:
    if (r > half) return q + 1;
    if (r == half) return q + (q & 1);
    return q;
  case RSPO_TRIM_ROUND_UP: return (v + mask) >> shift;
  case RSPO_TRIM_ROUND_TOWARDS_ZERO: return v >= 0 ? q : -((-v) >> shift);
  case RSPO_TRIM_ROUND_TOWARDS_INFINITY: return v >= 0 ? ((v + mask) >> shift) : -((-v + mask) >> shift);
  case RSPO_TRIM_ROUND_HALF_DOWN: return (v + half - 1) >> shift;
  case RSPO_TRIM_ROUND_HALF_TOWARDS_ZERO:
    return v >= 0 ? ((v + half - 1) >> shift) : -((-v + half - 1) >> shift);
  case RSPO_TRIM_ROUND_HALF_TOWARDS_INFINITY: return v >= 0 ? ((v + half) >> shift) : -((-v + half) >> shift);
  case RSPO_TRIM_ROUND_HALF_TO_ODD:
    if (r > half) return q + 1;
    if (r == half) return q + (1 - (q & 1));
    return q;
  default: return q;
  }
}

/* FixedPoint.fromDouble: math.round(x * 2^bp) = floor(x*2^bp + 0.5) (S9) */
static int64_t from_double(double x, int bp) { return (int64_t)floor(ldexp(x, bp) + 0.5); }

static int bitlen_u(uint32_t v) { int n = 0; while (v) { ++n; v >>= 1; } return n; }
static int ilog2_ceil(uint32_t v) { int n = 0; while ((1u << n) < v) ++n; return n; }
static int ilog2_floor(uint64_t v) { int n = 0; while (v >>= 1) ++n; return n; } /* chisel3.util.Log2, 0 for v<=1 (S10) */

static uint32_t bitrev(uint32_t v, int bits) {
  uint32_t r = 0;
  for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1u) << (bits - 1 - i);
  return r;
}

/* ------------------------------------------------------------------------------------------------
 * tables
 * ---------------------------------------------------------------------------------------------- */

/* SDFChainRadix22.scala:255-267: twiddle(t) = (cos, -sin)(2*pi*t/stageN) via fromDoubleWithFixedWidth into Q2.(tw-2).
 * The argument is evaluated as the reference writes it: 2 * Pi * t / stageN (left to right). */
void rspo_twiddle_table(int stage_n, int tw_width, int32_t *re, int32_t *im) {
  const int bp = tw_width - 2;
  const int n = 3 * stage_n / 4;
  for (int t = 0; t < n; ++t) {
    const double ang = 2.0 * M_PI * (double)t / (double)stage_n;
    re[t] = (int32_t)rspo_wrap(from_double(cos(ang), bp), tw_width);
    im[t] = (int32_t)rspo_wrap(from_double(-sin(ang), bp), tw_width);
  }
}

/* NCOTable.scala:254-258: T[i] = fromDoubleWithFixedWidth(sin(i * 0.5 * Pi / tableSize)) in Q2.(W-2) */
void rspo_nco_table(int table_size, int table_width, int32_t *tab) {
  for (int i = 0; i < table_size; ++i) {
    const double t = ((double)i * 0.5 * M_PI) / (double)table_size;
    tab[i] = (int32_t)from_double(sin(t), table_width - 2);
  }
}

/* LogMagMux.scala:120-131: LUT[n] = fromDoubleWithFixedWidth(log2(n/num + 1)) in FixedPoint((bposLog+1).W, bposLog.BP) */
void rspo_log2_table(int lookup_width, int bp_log, int32_t *tab) {
  const int num = 1 << lookup_width;
  const double ln2 = log(2.0);
  for (int n = 0; n < num; ++n) {
    const double v = log((double)n / (double)num + 1.0) / ln2;
    tab[n] = (int32_t)rspo_wrap(from_double(v, bp_log), bp_log + 1);
  }
}

/* ------------------------------------------------------------------------------------------------
 * FFT  (SDFChainRadix22.scala, SDFFFTUtil.scala:51-61)
 * ---------------------------------------------------------------------------------------------- */

/* DspComplex context_* then trimBinary(bpos), SDFChainRadix22.scala:298-308 (S5, S4, S6).
 * a: data (bp_a fractional bits), b: twiddle (bp_b).  Result wrapped to data_width. */
static void cmul_fixed(int64_t ar, int64_t ai, int64_t br, int64_t bi, int bp_a, int bp_b, int trim, int use4muls,
                       int data_width, int64_t *ore, int64_t *oim) {
  const int prod_bp = bp_a + bp_b;
  const int ctx_bp = (bp_a > bp_b ? bp_a : bp_b) + 1; /* binaryPointGrowth = 1 (DspContext default) */
  const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
  const int cur_bp = prod_bp - s1;
  int64_t re, im;
  if (use4muls) {
    re = rspo_trim(ar * br, s1, trim) - rspo_trim(ai * bi, s1, trim);
    im = rspo_trim(ar * bi, s1, trim) + rspo_trim(ai * br, s1, trim);
  } else {
    /* 3-multiplier structure of dsptools DspComplex.timesContext: every add/sub Grow, each product trimmed */
    const int64_t c_p_d = br + bi;
    const int64_t a_p_b = ar + ai;
    const int64_t b_m_a = ai - ar;
    const int64_t ac_p_ad = rspo_trim(ar * c_p_d, s1, trim);
    const int64_t ad_p_bd = rspo_trim(a_p_b * bi, s1, trim);
    const int64_t bc_m_ac = rspo_trim(b_m_a * br, s1, trim);
    re = ac_p_ad - ad_p_bd;
    im = ac_p_ad + bc_m_ac;
  }
  /* mulres.trimBinary(bpos): only trims when the current binary point is larger */
  if (cur_bp > bp_a) {
    re = rspo_trim(re, cur_bp - bp_a, trim);
    im = rspo_trim(im, cur_bp - bp_a, trim);
  } else if (cur_bp < bp_a) { /* `:=` aligns binary points */
    re <<= (bp_a - cur_bp);
    im <<= (bp_a - cur_bp);
  }
  *ore = rspo_wrap(re, data_width);
  *oim = rspo_wrap(im, data_width);
}

/* twiddle index pattern [0]*q ++ [2k] ++ [k] ++ [3k], SDFChainRadix22.scala:255-259 */
static inline int tw_index(int n, int q) {
  const int quarter = n / q, k = n % q;
  switch (quarter) { case 0: return 0; case 1: return 2 * k; case 2: return k; default: return 3 * k; }
}

/* twiddle tables of every multiplier stage of an S-stage transform (odd stage index i < S-1) */
typedef struct { int32_t *re[32]; int32_t *im[32]; } fft_tabs;
static void fft_tabs_free(fft_tabs *t) { for (int i = 0; i < 32; ++i) { free(t->re[i]); free(t->im[i]); t->re[i] = t->im[i] = NULL; } }
/* radix-2 stage ROM: twiddles_rom(k) = (cos, -sin)(2*Pi*k / (2*delay)), k < delay (SDFChainRadix2.scala:300-306) */
void rspo_twiddle_table_radix2(int delay, int tw_width, int32_t *re, int32_t *im) {
  const int bp = tw_width - 2;
  for (int k = 0; k < delay; ++k) {
    const double ang = 2.0 * M_PI * (double)k / (double)(2 * delay);
    re[k] = (int32_t)rspo_wrap(from_double(cos(ang), bp), tw_width);
    im[k] = (int32_t)rspo_wrap(from_double(-sin(ang), bp), tw_width);
  }
}

static int fft_tabs_build(fft_tabs *t, const rspo_fft_cfg *cfg) {
  const int S = cfg->log_n, tw = cfg->twiddle_width;
  memset(t, 0, sizeof(*t));
  if (cfg->sdf_radix == 2) { /* one ROM per stage, slot = log2(delay) */
    for (int d = 0; d < S; ++d) {
      const int delay = 1 << d;
      t->re[d] = (int32_t *)malloc(sizeof(int32_t) * (size_t)delay);
      t->im[d] = (int32_t *)malloc(sizeof(int32_t) * (size_t)delay);
      if (!t->re[d] || !t->im[d]) { fft_tabs_free(t); return -2; }
      rspo_twiddle_table_radix2(delay, tw, t->re[d], t->im[d]);
    }
    return 0;
  }
  if (cfg->decim_dit) { /* DIT: stageN = 2 << (index + 1) (SDFChainRadix22.scala:228); slot = index */
    for (int i = 1; i < S; ++i) {
      const int stage_n = 4 << i;
      t->re[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)stage_n);
      t->im[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)stage_n);
      if (!t->re[i] || !t->im[i]) { fft_tabs_free(t); return -2; }
      rspo_twiddle_table(stage_n, tw, t->re[i], t->im[i]);
    }
    return 0;
  }
  for (int i = 1; i < S - 1; i += 2) {
    const int stage_n = 4 << (S - 1 - i); /* 2 << (numStages - index) (SDFChainRadix22.scala:228) */
    t->re[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)stage_n);
    t->im[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)stage_n);
    if (!t->re[i] || !t->im[i]) { fft_tabs_free(t); return -2; }
    rspo_twiddle_table(stage_n, tw, t->re[i], t->im[i]);
  }
  return 0;
}

/* radix-2 SDF chain, DIF or DIT (SDFChainRadix2.scala:18-440).
 * Stage with delay D: Butterfly + div2 / keep-LSB exactly as the radix-2^2 stage (:398-440).  The twiddle multiplier sits
 * behind the butterfly on the difference output in DIF (`cntr < delay && cntr =/= 0`, :343-357: index k = 1..D-1 of the
 * second half) and in front of the butterfly on the second-half input in DIT (`cntr > delay`, :321-339).  The product is
 * taken with trimType = NoTrim / overflowType = Grow, i.e. exactly, and trimmed ONCE by trimBinary(bpos) under
 * params.trimType (:350-351), then assigned to protoIQ (MSBs dropped).  k = 0 bypasses the multiplier.
 * DIF: delays N/2 .. 1, natural-order input, bit-reversed output; DIT: delays 1 .. N/2, bit-reversed input, natural
 * output (:26, SDFFFT.scala:217-270: with useBitReverse the ping-pong buffer sits behind DIF / in front of DIT).
 * A run-time size of k stages is the last k stages in DIF and the first k in DIT (:127-134,196-203); the caller's
 * keep_msb_mask is already aligned to the stages of THIS transform. */
static inline int bf_too_big(int64_t v, int w_in);
static int fft_fixed_core_radix2(const rspo_fft_cfg *cfg, const fft_tabs *tabs, int64_t *xr, int64_t *xi,
                                 const int64_t *in_re, const int64_t *in_im, int64_t *out_re, int64_t *out_im, uint32_t *ovf) {
  const int S = cfg->log_n, N = 1 << S;
  const int w = cfg->data_width, tbp = cfg->twiddle_width - 2, trim = cfg->trim_type;
  const int dit = cfg->decim_dit;
  for (int p = 0; p < N; ++p) {
    const int src = (dit && cfg->bit_reverse) ? (int)bitrev((uint32_t)p, S) : p;
    xr[p] = rspo_wrap(cfg->fft_dir ? in_re[src] : in_im[src], w); /* ifft: swap on the way in (:214-220) */
    xi[p] = rspo_wrap(cfg->fft_dir ? in_im[src] : in_re[src], w);
  }
  for (int i = 0; i < S; ++i) {
    const int dl = dit ? i : S - 1 - i, D = 1 << dl;
    const int scale = (cfg->keep_msb_mask >> i) & 1;
    const int32_t *tr = tabs->re[dl], *ti = tabs->im[dl];
    for (int b = 0; b < N; b += 2 * D)
      for (int k = 0; k < D; ++k) {
        const int p0 = b + k, p1 = b + k + D;
        if (dit && k) {
          const int64_t ar = xr[p1], ai = xi[p1];
          xr[p1] = rspo_wrap(rspo_trim(ar * tr[k] - ai * ti[k], tbp, trim), w);
          xi[p1] = rspo_wrap(rspo_trim(ar * ti[k] + ai * tr[k], tbp, trim), w);
        }
        const int64_t sr = xr[p0] + xr[p1], si = xi[p0] + xi[p1];
        const int64_t dr = xr[p0] - xr[p1], di = xi[p0] - xi[p1];
        if (ovf && (bf_too_big(sr, w) || bf_too_big(si, w) || bf_too_big(dr, w) || bf_too_big(di, w))) *ovf |= 1u << i;
        if (scale) {
          xr[p0] = rspo_wrap(rspo_trim(sr, 1, trim), w); xi[p0] = rspo_wrap(rspo_trim(si, 1, trim), w);
          xr[p1] = rspo_wrap(rspo_trim(dr, 1, trim), w); xi[p1] = rspo_wrap(rspo_trim(di, 1, trim), w);
        } else {
          xr[p0] = rspo_wrap(sr, w); xi[p0] = rspo_wrap(si, w);
          xr[p1] = rspo_wrap(dr, w); xi[p1] = rspo_wrap(di, w);
        }
        if (!dit && k) {
          const int64_t ar = xr[p1], ai = xi[p1];
          xr[p1] = rspo_wrap(rspo_trim(ar * tr[k] - ai * ti[k], tbp, trim), w);
          xi[p1] = rspo_wrap(rspo_trim(ar * ti[k] + ai * tr[k], tbp, trim), w);
        }
      }
  }
  for (int n = 0; n < N; ++n) {
    const int p = (!dit && cfg->bit_reverse) ? (int)bitrev((uint32_t)n, S) : n;
    out_re[n] = (cfg->fft_dir ? xr[p] : xi[p]); /* ifft: swap on the way out, scalar 1.0 (:280,289-300) */
    out_im[n] = (cfg->fft_dir ? xi[p] : xr[p]);
  }
  return 0;
}

/* radix-2^2 SDF chain, DIT (SDFChainRadix22.scala:27,224-352 with decimType = DITDecimType).
 * Delays 1, 2, .. N/2 (:27), bit-reversed input, natural output (SDFFFT.scala:140-170: with useBitReverse the ping-pong buffer
 * sits in front).  What precedes the butterfly of stage `index` (delay D = 2^index) is decided by position alone:
 *   index 0: nothing (doNothing, :227);
 *   isMulTw (:229: index + 1 even when numStages is odd, odd when numStages is even -- numStages is the COMPILE-TIME stage count):
 *     the input at position p is multiplied by twiddle [0]*D ++ [2k] ++ [k] ++ [3k] (p mod 4D), stageN = 4D (:228,255-259,282-284),
 *     `stg_in context_* twiddles` under params.trimType, then trimBinary(bpos) (:298-308,314-316);
 *   otherwise: x(-j) on positions [3D/2, 2D) of every 2D block (:323,339-342).
 * Butterfly / div2 / keep-LSB as in every stage (:459-493).  A run-time size of k stages is the first k stages (:145-147,216-217). */
static int fft_fixed_core_r22_dit(const rspo_fft_cfg *cfg, const fft_tabs *tabs, int64_t *xr, int64_t *xi,
                                  const int64_t *in_re, const int64_t *in_im, int64_t *out_re, int64_t *out_im, uint32_t *ovf) {
  const int S = cfg->log_n, N = 1 << S;
  const int w = cfg->data_width, bp = cfg->bin_point, tw = cfg->twiddle_width, trim = cfg->trim_type;
  const int smax = cfg->num_stages > 0 ? cfg->num_stages : S;
  for (int p = 0; p < N; ++p) {
    const int src = cfg->bit_reverse ? (int)bitrev((uint32_t)p, S) : p;
    xr[p] = rspo_wrap(cfg->fft_dir ? in_re[src] : in_im[src], w);
    xi[p] = rspo_wrap(cfg->fft_dir ? in_im[src] : in_re[src], w);
  }
  for (int i = 0; i < S; ++i) {
    const int D = 1 << i;
    const int scale = (cfg->keep_msb_mask >> i) & 1;
    const int is_tw = i > 0 && (((i + 1) % 2 == 0 && smax % 2 == 1) || ((i + 1) % 2 == 1 && smax % 2 == 0));
    if (is_tw) {
      const int stage_n = 4 * D;
      const int32_t *tr = tabs->re[i], *ti = tabs->im[i];
      for (int p = 0; p < N; ++p) {
        const int t = tw_index(p % stage_n, D);
        cmul_fixed(xr[p], xi[p], tr[t], ti[t], bp, tw - 2, trim, cfg->use4muls, w, &xr[p], &xi[p]);
      }
    } else if (i > 0) {
      for (int b = 0; b < N; b += 2 * D)
        for (int k = 3 * D / 2; k < 2 * D; ++k) {
          const int p = b + k;
          const int64_t r = xr[p], im = xi[p];
          xr[p] = im;
          xi[p] = rspo_wrap(-r, w);
        }
    }
    for (int b = 0; b < N; b += 2 * D)
      for (int k = 0; k < D; ++k) {
        const int p0 = b + k, p1 = b + k + D;
        const int64_t sr = xr[p0] + xr[p1], si = xi[p0] + xi[p1];
        const int64_t dr = xr[p0] - xr[p1], di = xi[p0] - xi[p1];
        if (ovf && (bf_too_big(sr, w) || bf_too_big(si, w) || bf_too_big(dr, w) || bf_too_big(di, w))) *ovf |= 1u << i;
        if (scale) {
          xr[p0] = rspo_wrap(rspo_trim(sr, 1, trim), w); xi[p0] = rspo_wrap(rspo_trim(si, 1, trim), w);
          xr[p1] = rspo_wrap(rspo_trim(dr, 1, trim), w); xi[p1] = rspo_wrap(rspo_trim(di, 1, trim), w);
        } else {
          xr[p0] = rspo_wrap(sr, w); xi[p0] = rspo_wrap(si, w);
          xr[p1] = rspo_wrap(dr, w); xi[p1] = rspo_wrap(di, w);
        }
      }
  }
  for (int n = 0; n < N; ++n) {
    out_re[n] = (int32_t)(cfg->fft_dir ? xr[n] : xi[n]);
    out_im[n] = (int32_t)(cfg->fft_dir ? xi[n] : xr[n]);
  }
  return 0;
}

/* `tooBig` of the overflow check (SDFChainRadix22.scala:473-481) on a butterfly output grown from w_in to w_in + 1 bits */
static inline int bf_too_big(int64_t v, int w_in) { return v >= ((int64_t)1 << (w_in - 1)); }
#define BF_OVERFLOW(sr, si, dr, di, w_in) (bf_too_big(sr, w_in) || bf_too_big(si, w_in) || bf_too_big(dr, w_in) || bf_too_big(di, w_in))

int rspo_fft_stage_width(const rspo_fft_cfg *cfg, int stage) {
  int w = cfg->data_width + cfg->base_growth;
  for (int i = 0; i <= stage; ++i) w += (cfg->expand_mask >> i) & 1;
  return w;
}

static int fft_fixed_core(const rspo_fft_cfg *cfg, const fft_tabs *tabs, int64_t *xr, int64_t *xi,
                          const int64_t *in_re, const int64_t *in_im, int64_t *out_re, int64_t *out_im, uint32_t *ovf) {
  const int S = cfg->log_n;
  const int N = 1 << S;
  const int w0 = cfg->data_width, bp = cfg->bin_point, tw = cfg->twiddle_width, trim = cfg->trim_type;
  if (ovf) *ovf = 0;
  if ((cfg->expand_mask || cfg->base_growth) && (cfg->sdf_radix == 2 || cfg->decim_dit)) return -3; /* grow: radix 2^2 DIF only here */
  if (cfg->sdf_radix == 2) return fft_fixed_core_radix2(cfg, tabs, xr, xi, in_re, in_im, out_re, out_im, ovf);
  if (cfg->decim_dit) return fft_fixed_core_r22_dit(cfg, tabs, xr, xi, in_re, in_im, out_re, out_im, ovf);
  for (int p = 0; p < N; ++p) {
    /* ifft swaps real and imaginary on the way in (SDFChainRadix22.scala:205-211) */
    xr[p] = rspo_wrap(cfg->fft_dir ? in_re[p] : in_im[p], w0);
    xi[p] = rspo_wrap(cfg->fft_dir ? in_im[p] : in_re[p], w0);
  }
  for (int i = 0; i < S; ++i) {
    const int D = 1 << (S - 1 - i); /* SDFChainRadix22.scala:27,33 */
    const int scale = (cfg->keep_msb_mask >> i) & 1;
    const int grow = (cfg->expand_mask >> i) & 1;            /* useGrow (:156,453-457) */
    const int w = rspo_fft_stage_width(cfg, i);              /* protoIQstages(i) (SDFFFTParams.scala:119-123) */
    const int w_in = w - grow;                               /* the type the butterfly's inputs arrive in */
    for (int b = 0; b < N; b += 2 * D)
      for (int k = 0; k < D; ++k) {
        const int p0 = b + k, p1 = b + k + D;
        /* Butterfly: context_+ / context_- with Grow (SDFFFTUtil.scala:54-59) */
        const int64_t sr = xr[p0] + xr[p1], si = xi[p0] + xi[p1];
        const int64_t dr = xr[p0] - xr[p1], di = xi[p0] - xi[p1];
        if (grow) { /* the grown outputs are kept as they are; no overflow possible (:454-457) */
          xr[p0] = sr; xi[p0] = si; xr[p1] = dr; xi[p1] = di;
          continue;
        }
        if (ovf && BF_OVERFLOW(sr, si, dr, di, w_in)) *ovf |= 1u << i;
        if (scale) { /* div2(1), binaryPointGrowth 0, trimType (SDFChainRadix22.scala:459-464), stored to protoIQ (:492-493) */
          xr[p0] = rspo_wrap(rspo_trim(sr, 1, trim), w); xi[p0] = rspo_wrap(rspo_trim(si, 1, trim), w);
          xr[p1] = rspo_wrap(rspo_trim(dr, 1, trim), w); xi[p1] = rspo_wrap(rspo_trim(di, 1, trim), w);
        } else { /* keepLSB: the grown bit is dropped (:487-490) */
          xr[p0] = rspo_wrap(sr, w); xi[p0] = rspo_wrap(si, w);
          xr[p1] = rspo_wrap(dr, w); xi[p1] = rspo_wrap(di, w);
        }
      }
    if (i == S - 1) break; /* doNothing (:227) */
    if ((i & 1) == 0) {
      /* trivial -j on cntr in [D/2, D) of the feedback half, i.e. block positions [3D/2, 2D) (:322,327-329) */
      for (int b = 0; b < N; b += 2 * D)
        for (int k = 3 * D / 2; k < 2 * D; ++k) {
          const int p = b + k;
          const int64_t r = xr[p], im = xi[p];
          xr[p] = im;
          xi[p] = rspo_wrap(-r, w);
        }
    } else {
      const int stage_n = 4 * D; /* 2 << (numStages - index) (:228) */
      const int q = stage_n / 4;
      const int32_t *tr = tabs->re[i], *ti = tabs->im[i];
      for (int p = 0; p < N; ++p) {
        const int t = tw_index(p % stage_n, q);
        cmul_fixed(xr[p], xi[p], tr[t], ti[t], bp, tw - 2, trim, cfg->use4muls, w, &xr[p], &xi[p]);
      }
    }
  }
  for (int n = 0; n < N; ++n) {
    /* DIF output is bit-reversed; BitReversePingPong reads at Reverse(cnt) (BitReversePingPong.scala:93-109) */
    const int p = cfg->bit_reverse ? (int)bitrev((uint32_t)n, S) : n;
    /* ifft swaps again on the way out; scalar is 1.0 when no stage grows (:362,366-372) */
    out_re[n] = (cfg->fft_dir ? xr[p] : xi[p]);
    out_im[n] = (cfg->fft_dir ? xi[p] : xr[p]);
  }
  return 0;
}

int rspo_fft_fixed(const rspo_fft_cfg *cfg, const int32_t *in_re, const int32_t *in_im, int32_t *out_re,
                   int32_t *out_im) {
  return rspo_fft_fixed_ex(cfg, in_re, in_im, out_re, out_im, NULL);
}

int rspo_fft_fixed_ex(const rspo_fft_cfg *cfg, const int32_t *in_re, const int32_t *in_im, int32_t *out_re, int32_t *out_im,
                      uint32_t *overflow_mask) {
  const int S = cfg->log_n;
  if (S < 1 || S > 24 || cfg->data_width < 2 || cfg->data_width > 32) return -1;
  if (rspo_fft_stage_width(cfg, S - 1) > 32) return -1; /* int32 interface; wider results through rspo_fft_fixed_words64 */
  const int N = 1 << S;
  fft_tabs tabs;
  if (fft_tabs_build(&tabs, cfg)) return -2;
  int64_t *xr = (int64_t *)malloc(sizeof(int64_t) * 6 * (size_t)N);
  if (!xr) { fft_tabs_free(&tabs); return -2; }
  int64_t *ir = xr + 2 * N, *ii = xr + 3 * N, *orr = xr + 4 * N, *oi = xr + 5 * N;
  for (int p = 0; p < N; ++p) { ir[p] = in_re[p]; ii[p] = in_im[p]; }
  const int rc = fft_fixed_core(cfg, &tabs, xr, xr + N, ir, ii, orr, oi, overflow_mask);
  for (int p = 0; p < N; ++p) { out_re[p] = (int32_t)orr[p]; out_im[p] = (int32_t)oi[p]; }
  free(xr);
  fft_tabs_free(&tabs);
  return rc;
}

int rspo_fft_f64(int log_n, int bit_reverse, const double *in_re, const double *in_im, double *out_re,
                 double *out_im) {
  const int S = log_n, N = 1 << S;
  double *xr = (double *)malloc(sizeof(double) * N), *xi = (double *)malloc(sizeof(double) * N);
  if (!xr || !xi) { free(xr); free(xi); return -2; }
  memcpy(xr, in_re, sizeof(double) * N);
  memcpy(xi, in_im, sizeof(double) * N);
  for (int i = 0; i < S; ++i) {
    const int D = 1 << (S - 1 - i);
    for (int b = 0; b < N; b += 2 * D)
      for (int k = 0; k < D; ++k) {
        const int p0 = b + k, p1 = b + k + D;
        const double sr = xr[p0] + xr[p1], si = xi[p0] + xi[p1], dr = xr[p0] - xr[p1], di = xi[p0] - xi[p1];
        xr[p0] = 0.5 * sr; xi[p0] = 0.5 * si; xr[p1] = 0.5 * dr; xi[p1] = 0.5 * di;
      }
    if (i == S - 1) break;
    if ((i & 1) == 0) {
      for (int b = 0; b < N; b += 2 * D)
        for (int k = 3 * D / 2; k < 2 * D; ++k) {
          const int p = b + k; const double r = xr[p];
          xr[p] = xi[p]; xi[p] = -r;
        }
    } else {
      const int stage_n = 4 * D, q = stage_n / 4;
      for (int p = 0; p < N; ++p) {
        const int t = tw_index(p % stage_n, q);
        const double ang = 2.0 * M_PI * (double)t / (double)stage_n, c = cos(ang), s = -sin(ang);
        const double r = xr[p] * c - xi[p] * s, im = xr[p] * s + xi[p] * c;
        xr[p] = r; xi[p] = im;
      }
    }
  }
  for (int n = 0; n < N; ++n) {
    const int p = bit_reverse ? (int)bitrev((uint32_t)n, S) : n;
    out_re[n] = xr[p]; out_im[n] = xi[p];
  }
  free(xr); free(xi);
  return 0;
}

/* beat = Cat(re, im) with dataWidth-bit halves (FFTBlock.scala:106,114; FFTt/TesterUtil.scala:204-208) */
static inline void unpack_iq(uint32_t wd, int w, int32_t *re, int32_t *im) {
  *re = (int32_t)rspo_wrap((int64_t)(wd >> 16), w);
  *im = (int32_t)rspo_wrap((int64_t)(wd & 0xFFFFu), w);
}
static inline uint32_t pack_iq(int32_t re, int32_t im) {
  return (((uint32_t)re & 0xFFFFu) << 16) | ((uint32_t)im & 0xFFFFu);
}

static int fft_fixed_words_tabs(const rspo_fft_cfg *cfg, const fft_tabs *tabs, const uint32_t *in, uint32_t *out,
                                int n_chirps) {
  const int N = 1 << cfg->log_n;
  int rc = 0;
  int64_t *buf = (int64_t *)malloc(sizeof(int64_t) * 4 * (size_t)N);
  int64_t *x = (int64_t *)malloc(sizeof(int64_t) * 2 * (size_t)N);
  if (!buf || !x) { free(buf); free(x); return -2; }
  int64_t *ir = buf, *ii = buf + N, *orr = buf + 2 * N, *oi = buf + 3 * N;
  for (int c = 0; c < n_chirps && rc == 0; ++c) {
    for (int p = 0; p < N; ++p) { int32_t a, b; unpack_iq(in[(size_t)c * N + p], cfg->data_width, &a, &b); ir[p] = a; ii[p] = b; }
    rc = fft_fixed_core(cfg, tabs, x, x + N, ir, ii, orr, oi, NULL);
    for (int p = 0; p < N; ++p) out[(size_t)c * N + p] = pack_iq((int32_t)orr[p], (int32_t)oi[p]);
  }
  free(buf); free(x);
  return rc;
}

/* AXI4FFTBlock with wide data or grow stages: 64-bit beats (beatBytes = 8), or 128-bit beats (two uint64 words per beat, low word
 * first) when the output Cat(re, im) needs more than 64 bits.  in = low 2 * dataWidth bits Cat(re, im) of a 64-bit beat
 * (`in.bits.data.asTypeOf(params.protoIQ)`, FFTBlock.scala:106), out = Cat(re, im) of the last stage's type
 * (`fft.io.out.bits.asUInt`, :113; SDFFFT.scala:20).  overflow_mask (may be NULL): OR over the chirps of rspo_fft_fixed_ex's mask. */
int rspo_fft_fixed_words64(const rspo_fft_cfg *cfg, const uint64_t *in, uint64_t *out, int n_chirps, uint32_t *overflow_mask) {
  const int S = cfg->log_n, N = 1 << S, w0 = cfg->data_width;
  if (S < 1 || S > 24 || w0 < 2 || w0 > 32) return -1;
  const int wo = rspo_fft_stage_width(cfg, S - 1);
  if (wo > 40) return -1;
  const int words = 2 * wo > 64 ? 2 : 1;
  fft_tabs tabs;
  if (fft_tabs_build(&tabs, cfg)) return -2;
  int64_t *buf = (int64_t *)malloc(sizeof(int64_t) * 6 * (size_t)N);
  if (!buf) { fft_tabs_free(&tabs); return -2; }
  int64_t *x = buf, *ir = buf + 2 * N, *ii = buf + 3 * N, *orr = buf + 4 * N, *oi = buf + 5 * N;
  int rc = 0;
  uint32_t all = 0;
  const uint64_t mo = (1ULL << wo) - 1;
  for (int c = 0; c < n_chirps && rc == 0; ++c) {
    for (int p = 0; p < N; ++p) {
      const uint64_t wd = in[(size_t)c * N + p];
      ir[p] = rspo_wrap((int64_t)(wd >> w0), w0);
      ii[p] = rspo_wrap((int64_t)wd, w0);
    }
    uint32_t m = 0;
    rc = fft_fixed_core(cfg, &tabs, x, x + N, ir, ii, orr, oi, &m);
    all |= m;
    for (int p = 0; p < N; ++p) {
      const uint64_t re = (uint64_t)orr[p] & mo, im = (uint64_t)oi[p] & mo;
      uint64_t *o = out + ((size_t)c * N + p) * (size_t)words;
      o[0] = (wo < 64 ? (re << wo) : 0) | im;          /* bits 63..0 of Cat(re, im) */
      if (words == 2) o[1] = re >> (64 - wo);          /* bits 127..64 */
    }
  }
  if (overflow_mask) *overflow_mask = all;
  free(buf); fft_tabs_free(&tabs);
  return rc;
}

/* 32-bit beats, plus the overflow mask OR-ed over the chirps */
int rspo_fft_fixed_words_ovf(const rspo_fft_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps, uint32_t *overflow_mask) {
  const int N = 1 << cfg->log_n;
  if (cfg->data_width > 16 || cfg->log_n < 1 || cfg->log_n > 24 || cfg->expand_mask || cfg->base_growth) return -1;
  fft_tabs tabs;
  if (fft_tabs_build(&tabs, cfg)) return -2;
  int64_t *buf = (int64_t *)malloc(sizeof(int64_t) * 6 * (size_t)N);
  if (!buf) { fft_tabs_free(&tabs); return -2; }
  int64_t *x = buf, *ir = buf + 2 * N, *ii = buf + 3 * N, *orr = buf + 4 * N, *oi = buf + 5 * N;
  int rc = 0;
  uint32_t all = 0;
  for (int c = 0; c < n_chirps && rc == 0; ++c) {
    for (int p = 0; p < N; ++p) { int32_t a, b; unpack_iq(in[(size_t)c * N + p], cfg->data_width, &a, &b); ir[p] = a; ii[p] = b; }
    uint32_t m = 0;
    rc = fft_fixed_core(cfg, &tabs, x, x + N, ir, ii, orr, oi, &m);
    all |= m;
    for (int p = 0; p < N; ++p) out[(size_t)c * N + p] = pack_iq((int32_t)orr[p], (int32_t)oi[p]);
  }
  if (overflow_mask) *overflow_mask = all;
  free(buf); fft_tabs_free(&tabs);
  return rc;
}

int rspo_fft_fixed_words(const rspo_fft_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps) {
  if (cfg->data_width > 16 || cfg->log_n < 1 || cfg->log_n > 24) return -1;
  fft_tabs tabs;
  if (fft_tabs_build(&tabs, cfg)) return -2;
  const int rc = fft_fixed_words_tabs(cfg, &tabs, in, out, n_chirps);
  fft_tabs_free(&tabs);
  return rc;
}

/* ------------------------------------------------------------------------------------------------
 * LogMagMux (LogMagMux.scala:92-169)
 * ---------------------------------------------------------------------------------------------- */

int32_t rspo_mag_one(const rspo_mag_cfg *cfg, const int32_t *log_tab, int32_t re, int32_t im, int32_t *out_re) {
  const int w = cfg->data_width, bp = cfg->bin_point;
  if (out_re) *out_re = 0; /* .real := 0 for sel 1..3 (:147-163) */
  if (cfg->sel == 0) { if (out_re) *out_re = re; return im; }
  /* Real[T].abs under overflowType Grow: -(-2^(w-1)) = +2^(w-1) is representable (S8; corner) */
  const int64_t ar = re < 0 ? -(int64_t)re : re, ai = im < 0 ? -(int64_t)im : im;
  if (cfg->sel == 1) {
    /* (|re| context_* |re|) context_+ (|im| context_* |im|): product BP 2bp trimmed to bp+1 under the context trim
     * type, Grow add, then `:=` into proto.real (floor to bp, wrap to w) (:107-108; S5, S6) */
    const int prod_bp = 2 * bp, ctx_bp = bp + 1;
    const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
    const int cur_bp = prod_bp - s1;
    int64_t v = rspo_trim(ar * ar, s1, cfg->ctx_trim) + rspo_trim(ai * ai, s1, cfg->ctx_trim);
    if (cur_bp > bp) v >>= (cur_bp - bp); else if (cur_bp < bp) v <<= (bp - cur_bp);
    return (int32_t)rspo_wrap(v, w);
  }
  const int64_t u = ar > ai ? ar : ai, v = ar > ai ? ai : ar;
  const int64_t op1 = u + (v >> 3);                 /* U + V/8 (:99-100) */
  const int64_t op2 = (u - (u >> 3)) + (v >> 1);    /* 7/8 U + V/2 (:101-104) */
  const int64_t jpl = op1 > op2 ? op1 : op2;        /* (:110) */
  if (cfg->sel == 2) return (int32_t)rspo_wrap(jpl, w); /* jplMagOut.imag := jplMag (:156) */
  /* sel 3: log2 magnitude, the intended per-sample arithmetic (:115-143); see DESIGN.md for the RTL pipeline skew */
  const int wj = w + 2; /* width of jplMag: two Grow adds on op2 */
  const uint64_t ju = (uint64_t)jpl & ((1ULL << wj) - 1);
  const int k = ilog2_floor(ju);                    /* Log2(jplMag.asUInt()) (:115) */
  const int wk_a = ilog2_ceil((uint32_t)wj);        /* width of Log2's result */
  int wk = wk_a;
  { int wb = bitlen_u((uint32_t)bp); if (wb < 1) wb = 1; if (wb > wk) wk = wb; }
  const int64_t log_sint = rspo_wrap((int64_t)((uint32_t)(k - bp) & ((1u << wk) - 1)), wk); /* (logUInt - bpos.U).asSInt (:116) */
  const uint64_t no_lead_one = (ju - (1ULL << k)) & ((1ULL << wj) - 1);                    /* (:133) */
  int ws = bitlen_u((uint32_t)w); if (wk_a > ws) ws = wk_a;
  const uint32_t shift_num = (uint32_t)(w - k - 1) & ((1u << ws) - 1);                    /* (:134) */
  const int L = cfg->log2_lookup_width;
  const uint64_t shifted = shift_num < 63 ? (no_lead_one << shift_num) : 0;
  const uint32_t addr = (uint32_t)((shifted >> (w - L - 1)) & ((1ULL << L) - 1));         /* bits (w-2, w-L-1) (:136) */
  const int64_t frac = log_tab[addr];
  const int64_t lg = rspo_wrap((log_sint << cfg->bin_point_log) + frac, cfg->data_width_log); /* (:141) */
  /* log2Mag.asTypeOf(proto.real): raw-bit reinterpretation (:161) */
  const uint64_t raw = (uint64_t)lg & ((cfg->data_width_log >= 64) ? ~0ULL : ((1ULL << cfg->data_width_log) - 1));
  return (int32_t)(cfg->data_width_log >= w ? rspo_wrap((int64_t)raw, w) : (int64_t)raw);
}

int rspo_mag_words(const rspo_mag_cfg *cfg, const uint32_t *in, uint32_t *out, int64_t n) {
  int32_t *tab = NULL;
  if (cfg->sel == 3) {
    tab = (int32_t *)malloc(sizeof(int32_t) << cfg->log2_lookup_width);
    if (!tab) return -2;
    rspo_log2_table(cfg->log2_lookup_width, cfg->bin_point_log, tab);
  }
  for (int64_t i = 0; i < n; ++i) {
    int32_t re, im, ore;
    unpack_iq(in[i], cfg->data_width, &re, &im);
    const int32_t v = rspo_mag_one(cfg, tab, re, im, &ore);
    out[i] = pack_iq(ore, v);
  }
  free(tab);
  return 0;
}

/* the closed form the reference's testers use (RspChainTesterUtils.scala:120-127, LogMagMuxSpec.scala:23-31) */
double rspo_jpl_f64(double re, double im) {
  const double a = fabs(re), b = fabs(im), u = a > b ? a : b, v = a > b ? b : a;
  const double x = u + v / 8.0, y = 7.0 * u / 8.0 + v / 2.0;
  return x > y ? x : y;
}

/* ------------------------------------------------------------------------------------------------
 * CFAR cores (CFARCoreWithMem.scala, CFARCoreWithLis.scala, CFARCoreWithASR.scala)
 * ---------------------------------------------------------------------------------------------- */

static int cmp_desc_i32(const void *a, const void *b) {
  const int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return (x < y) - (x > y);
}

/* sumT = (io.in.bits * log2Ceil(leadLaggWindowSize)).cloneType (CFARCoreWithMem.scala:34): input width plus the
 * width of the signed literal */
static int cfar_sum_width(const rspo_cfar_cfg *c) {
  return c->in_width + bitlen_u((uint32_t)ilog2_ceil((uint32_t)c->lead_lagg_window_size)) + 1;
}

static inline int64_t xat(const int32_t *x, int n, int i) { return (i >= 0 && i < n) ? x[i] : 0; }

int rspo_cfar_fixed(const rspo_cfar_cfg *c, const int32_t *x, int64_t *thr_full, int32_t *thr_out, uint8_t *peak) {
  const int N = c->fft_win, W = c->window_cells, G = c->guard_cells;
  /* run-time asserts of the cores (CFARCoreWithMem.scala:31-32) */
  if (!(N > 2 * W + 2 * G + 1) || W < 1 || G < 0) return -1;
  if (c->core == RSPO_CFAR_CORE_MEM && G < 1) return -1;
  const int sumw = cfar_sum_width(c);
  const int bp_in = c->in_bp, bp_s = c->scaler_bp;
  const int64_t scaler = rspo_wrap(c->threshold_scaler, c->scaler_width);
  int32_t *sorted = NULL;
  if (c->core == RSPO_CFAR_CORE_LIS || c->core == RSPO_CFAR_CORE_OS) { sorted = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)W); if (!sorted) return -2; }
  int rc = 0;
  for (int n = 0; n < N; ++n) {
    const int lagg_full = (n + G + W <= N - 1);
    const int lead_full = (n - G - W >= 0);
    int64_t L = 0, R = 0, thr0 = 0;
    if (c->core == RSPO_CFAR_CORE_MEM || (c->core == RSPO_CFAR_CORE_LIS && c->cfar_algorithm == 0)) {
      /* running sums of the full windows (CFARCoreWithMem.scala:102-130), registers of type sumT wrap (S11) */
      int64_t sl = 0, sr = 0;
      if (lagg_full) for (int j = n + G + 1; j <= n + G + W; ++j) sl += x[j];
      if (lead_full) for (int j = n - G - W; j <= n - G - 1; ++j) sr += x[j];
      L = rspo_wrap(sl, sumw) >> c->div_sum; /* BinaryRepresentation.shr: arithmetic (S7) (:132-133) */
      R = rspo_wrap(sr, sumw) >> c->div_sum;
    } else if (c->core == RSPO_CFAR_CORE_OS) {
      /* cfarOS (CFARUtils.scala:165-171): thr = (lead ++ lagg).sorted((1.5 * referenceCells).toInt), ascending; only where both
       * windows exist (the golden's loop runs ix = windowCells until totalCells - windowCells and pads the rest with 0) */
      if (lagg_full && lead_full) {
        for (int j = 0; j < W; ++j) { sorted[j] = x[n - G - W + j]; sorted[W + j] = x[n + G + 1 + j]; }
        qsort(sorted, 2 * (size_t)W, sizeof(int32_t), cmp_desc_i32);
        L = R = sorted[2 * W - 1 - (3 * W) / 2];   /* ascending index floor(1.5 W) of 2 W cells */
      }
    } else if (c->core == RSPO_CFAR_CORE_LIS) {
      /* descending sorted windows, 1-based rank from the largest (CFARCoreWithLis.scala:32-39,136-143) */
      if (c->index_lagg < 1 || c->index_lagg > W || c->index_lead < 1 || c->index_lead > W) { rc = -1; break; }
      if (lagg_full) { for (int j = 0; j < W; ++j) sorted[j] = x[n + G + 1 + j]; qsort(sorted, W, sizeof(int32_t), cmp_desc_i32); L = sorted[c->index_lagg - 1]; }
      if (lead_full) { for (int j = 0; j < W; ++j) sorted[j] = x[n - G - W + j]; qsort(sorted, W, sizeof(int32_t), cmp_desc_i32); R = sorted[c->index_lead - 1]; }
    } else { /* ASR core: sub-window 0 sums for CA/GO/SO (CFARCoreWithASR.scala:285-288) */
      const int sub = c->sub_cells;
      if (sub < 1 || sub > W || (sub & (sub - 1)) || (W % sub)) { rc = -1; break; }
      int64_t sl = 0, sr = 0;
      if (lagg_full) for (int j = 0; j < sub; ++j) sl += x[n + G + W - j];   /* newest `sub` cells of the lagging window */
      if (lead_full) for (int j = 0; j < sub; ++j) sr += x[n - G - 1 - j];   /* newest `sub` cells of the leading window */
      L = rspo_wrap(sl, sumw) >> c->div_sum;
      R = rspo_wrap(sr, sumw) >> c->div_sum;
    }
    const int64_t go = L > R ? L : R, so = L < R ? L : R;
    int64_t by_mode;
    switch (c->cfar_mode) { /* MuxLookup, default smallestOf (CFARCoreWithMem.scala:138-141) */
    case 0: by_mode = (R + L) >> 1; break;
    case 1: by_mode = go; break;
    case 2: by_mode = so; break;
    default: by_mode = so; break;
    }
    if (c->core == RSPO_CFAR_CORE_OS) by_mode = L;
    if (c->core == RSPO_CFAR_CORE_ASR && c->cfar_mode == 3 && lagg_full && lead_full) {
      /* CASH: min over sub-windows of max(laggSum_i, leadSum_i), >> log2(subCells) (CFARCoreWithASR.scala:193-209) */
      const int sub = c->sub_cells, ns = W / sub;
      int64_t mn = 0;
      for (int i = 0; i < ns; ++i) {
        int64_t sl = 0, sr = 0;
        for (int j = 0; j < sub; ++j) { sl += x[n + G + W - i * sub - j]; sr += x[n - G - 1 - i * sub - j]; }
        sl = rspo_wrap(sl, sumw); sr = rspo_wrap(sr, sumw);
        const int64_t mx = sl > sr ? sl : sr;
        if (i == 0 || mx < mn) mn = mx;
      }
      by_mode = mn >> ilog2_floor((uint64_t)sub);
    }
    if (c->core == RSPO_CFAR_CORE_MEM) {
      /* one-sided windows at the edges (CFARCoreWithMem.scala:154-171) */
      if (!lead_full && !lagg_full) thr0 = 0;
      else if (lagg_full && !lead_full) thr0 = L;
      else if (!lagg_full) thr0 = R; /* enableRightThr */
      else thr0 = by_mode;
    } else {
      /* threshold forced to 0 unless both windows are full (CFARCoreWithLis.scala:172-173, CFARCoreWithASR.scala:311-315) */
      thr0 = (lagg_full && lead_full) ? by_mode : 0;
    }
    /* scaling (CFARCoreWithMem.scala:173-177): context_* (BP max+1, context trim) or context_+ (Grow) */
    int64_t thr; int thr_bp;
    if (c->log_or_linear) {
      const int prod_bp = bp_in + bp_s, ctx_bp = (bp_in > bp_s ? bp_in : bp_s) + 1;
      const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
      thr = rspo_trim(thr0 * scaler, s1, c->ctx_trim);
      thr_bp = prod_bp - s1;
    } else {
      thr_bp = bp_in > bp_s ? bp_in : bp_s;
      thr = (thr0 << (thr_bp - bp_in)) + (scaler << (thr_bp - bp_s));
    }
    const int64_t cut = x[n];
    const int64_t cut_al = cut << (thr_bp - bp_in); /* compare at full precision (:185) */
    int is_peak = cut_al > thr;
    if (c->core == RSPO_CFAR_CORE_OS && !(lagg_full && lead_full)) { is_peak = 0; thr = 0; }   /* no edge bins in the golden's loop; padded with 0.0 */
    if (c->peak_grouping) {
      /* neighbours come from guard shift registers at their compile-time ends (parallelOut.last / .head);
       * registers past the run-time depth never load and read 0.  The Mem core takes BOTH from the lagging
       * guard (CFARCoreWithMem.scala:182-183); the Lis/ASR cores take the right one from the leading guard
       * (CFARCoreWithLis.scala:191-192, CFARCoreWithASR.scala:326-327). Out-of-chirp cells are defined as 0. */
      const int gfull = (G == c->guard_window_size);
      int64_t left, right;
      if (G == 0) { left = 0; right = 0; } /* the guardCells REGISTER may be 0 on these cores (CFARParams.scala:41 only bounds the compile-time guardWindowSize): the reference then substitutes the window ends (CFARCoreWithLis.scala:191-192, CFARCoreWithASR.scala:326-327); not restated -- the product refuses peakGrouping with guardCells = 0 on the Lis / ASR cores (RSP_EUNSUPPORTED) */
      else {
        left = gfull ? xat(x, N, n + 1) : 0;
        right = (c->core == RSPO_CFAR_CORE_MEM) ? xat(x, N, n + G) : xat(x, N, n - 1);
      }
      is_peak = is_peak && (cut > left) && (cut > right);
    }
    if (thr_full) thr_full[n] = thr;
    /* io.out.bits.threshold := threshold : re-shape to protoThreshold (floor LSBs, wrap MSBs) (S6) */
    int64_t t = thr;
    if (thr_bp > c->thr_bp) t >>= (thr_bp - c->thr_bp); else if (thr_bp < c->thr_bp) t <<= (c->thr_bp - thr_bp);
    if (thr_out) thr_out[n] = (int32_t)rspo_wrap(t, c->thr_width);
    if (peak) peak[n] = (uint8_t)is_peak;
  }
  free(sorted);
  return rc;
}

int rspo_cfar_words(const rspo_cfar_cfg *c, const uint32_t *in, uint32_t *out, int n_chirps) {
  const int N = c->fft_win;
  const int binw = ilog2_ceil((uint32_t)c->fft_size);
  int32_t *x = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)N);
  uint8_t *pk = (uint8_t *)malloc((size_t)N);
  if (!x || !pk) { free(x); free(pk); return -2; }
  int32_t *th = x + N;
  int rc = 0;
  for (int ch = 0; ch < n_chirps && rc == 0; ++ch) {
    /* in.bits.data.asTypeOf(protoIn): low in_width bits (CFARDspBlock.scala:49) */
    for (int n = 0; n < N; ++n) x[n] = (int32_t)rspo_wrap((int64_t)in[(size_t)ch * N + n], c->in_width);
    rc = rspo_cfar_fixed(c, x, NULL, th, pk);
    for (int n = 0; n < N; ++n) {
      uint64_t wd;
      if (c->send_cut) wd = (uint64_t)((uint32_t)x[n]) & ((1ULL << c->in_width) - 1); /* (:78-79) */
      else /* Cat(threshold, fftBin, peak) (:90) */
        wd = ((((uint64_t)(uint32_t)th[n]) & ((1ULL << c->thr_width) - 1)) << (binw + 1)) |
             (((uint64_t)n & ((1ULL << binw) - 1)) << 1) | (uint64_t)pk[n];
      out[(size_t)ch * N + n] = (uint32_t)wd;
    }
  }
  free(x); free(pk);
  return rc;
}

static int cmp_desc_f64_(const void *a, const void *b) {
  const double x = *(const double *)a, y = *(const double *)b;
  return (x < y) - (x > y);
}
/* The DspReal flavour of the three cores: the same window / edge / mode / peak-grouping policy as rspo_cfar_fixed above
 * (CFARCoreWithMem.scala:102-189, CFARCoreWithLis.scala:136-194, CFARCoreWithASR.scala:193-209,285-332) with doubles in place of
 * the fixed-point registers: no sum wrap, `>> divSum` is a division by 2^divSum, CASH divides by subCells. */
int rspo_cfar_f64(const rspo_cfar_cfg *c, double scaler, const double *x, double *thr, uint8_t *peak) {
  const int N = c->fft_win, W = c->window_cells, G = c->guard_cells;
  if (!(N > 2 * W + 2 * G + 1) || W < 1 || G < 0) return -1;
  if (c->core == RSPO_CFAR_CORE_MEM && G < 1) return -1;
  const double div = ldexp(1.0, c->div_sum);
  const int gos = (c->core == RSPO_CFAR_CORE_LIS && c->cfar_algorithm != 0);
  const int os = c->core == RSPO_CFAR_CORE_OS;
  double *sorted = NULL;
  if (gos || os) {
    if (gos && (c->index_lagg < 1 || c->index_lagg > W || c->index_lead < 1 || c->index_lead > W)) return -1;
    sorted = (double *)malloc(sizeof(double) * 2 * (size_t)W);
    if (!sorted) return -2;
  }
  const int sub = c->sub_cells;
  if (c->core == RSPO_CFAR_CORE_ASR && (sub < 1 || sub > W || (sub & (sub - 1)) || (W % sub))) return -1;
  for (int n = 0; n < N; ++n) {
    const int lagg_full = (n + G + W <= N - 1), lead_full = (n - G - W >= 0);
    double L = 0, R = 0;
    if (os) {
      if (lagg_full && lead_full) {
        for (int j = 0; j < W; ++j) { sorted[j] = x[n - G - W + j]; sorted[W + j] = x[n + G + 1 + j]; }
        qsort(sorted, 2 * (size_t)W, sizeof(double), cmp_desc_f64_);
        L = R = sorted[2 * W - 1 - (3 * W) / 2];
      }
    } else if (gos) {
      if (lagg_full) { for (int j = 0; j < W; ++j) sorted[j] = x[n + G + 1 + j]; qsort(sorted, (size_t)W, sizeof(double), cmp_desc_f64_); L = sorted[c->index_lagg - 1]; }
      if (lead_full) { for (int j = 0; j < W; ++j) sorted[j] = x[n - G - W + j]; qsort(sorted, (size_t)W, sizeof(double), cmp_desc_f64_); R = sorted[c->index_lead - 1]; }
    } else {
      const int wl = c->core == RSPO_CFAR_CORE_ASR ? sub : W;   /* ASR: the newest subCells cells of each window (:285-288) */
      double sl = 0, sr = 0;
      if (lagg_full) for (int j = 0; j < wl; ++j) sl += x[n + G + W - j];
      if (lead_full) for (int j = 0; j < wl; ++j) sr += x[n - G - 1 - j];
      L = sl / div; R = sr / div;
    }
    double by_mode = c->cfar_mode == 0 ? 0.5 * (L + R) : (c->cfar_mode == 1 ? (L > R ? L : R) : (L < R ? L : R));
    if (os) by_mode = L;
    if (c->core == RSPO_CFAR_CORE_ASR && c->cfar_mode == 3 && lagg_full && lead_full) {
      const int ns = W / sub;
      double mn = 0;
      for (int i = 0; i < ns; ++i) {
        double sl = 0, sr = 0;
        for (int j = 0; j < sub; ++j) { sl += x[n + G + W - i * sub - j]; sr += x[n - G - 1 - i * sub - j]; }
        const double mx = sl > sr ? sl : sr;
        if (i == 0 || mx < mn) mn = mx;
      }
      by_mode = mn / (double)sub;
    }
    double t0;
    if (c->core == RSPO_CFAR_CORE_MEM) {
      if (!lead_full && !lagg_full) t0 = 0; else if (!lead_full) t0 = L; else if (!lagg_full) t0 = R; else t0 = by_mode;
    } else t0 = (lagg_full && lead_full) ? by_mode : 0;
    double t = c->log_or_linear ? t0 * scaler : t0 + scaler;
    int pk = x[n] > t;
    if (os && !(lagg_full && lead_full)) { pk = 0; t = 0; }
    if (c->peak_grouping) {
      const int gfull = (G == c->guard_window_size);
      double left = 0, right = 0;
      if (G != 0) {
        left = (gfull && n + 1 < N) ? x[n + 1] : 0.0;
        if (c->core == RSPO_CFAR_CORE_MEM) right = (n + G < N) ? x[n + G] : 0.0;
        else right = (n - 1 >= 0) ? x[n - 1] : 0.0;
      }
      pk = pk && x[n] > left && x[n] > right;
    }
    thr[n] = t; peak[n] = (uint8_t)pk;
  }
  free(sorted);
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Scala golden models, restated (CFARUtils.scala)
 * ---------------------------------------------------------------------------------------------- */

/* Seq.slice(from, until).sum with Scala's clamping semantics */
static double slice_sum(const double *s, int n, int from, int until) {
  if (from < 0) from = 0;
  if (until > n) until = n;
  double a = 0.0;
  for (int i = from; i < until; ++i) a += s[i];
  return a;
}
static int cmp_asc_f64(const void *a, const void *b) {
  const double x = *(const double *)a, y = *(const double *)b;
  return (x > y) - (x < y);
}
static int cmp_desc_f64(const void *a, const void *b) { return -cmp_asc_f64(a, b); }

/* cfarCA (CFARUtils.scala:84-156) */
int rspo_golden_cfar_ca(const double *sig, int n, int mode, int R, int G, int consider_edges, double scaling,
                        int log_mode, double *thr, int *n_thr, int *peaks, int *n_peaks) {
  if (!(R > 0 && G > 0)) return -1; /* require (:85) */
  const int wc = R + G;
  int nt = 0, np = 0;
  if (consider_edges) { /* (:91-113) */
    for (int ix = 0; ix < n; ++ix) {
      const double lead = (ix < wc) ? 0.0 : slice_sum(sig, n, ix - wc, ix - G) / R;
      const double lagg = (ix >= n - wc) ? 0.0 : slice_sum(sig, n, ix + G + 1, ix + wc + 1) / R;
      double t;
      if (ix < wc) t = lagg;
      else if (ix >= n - wc) t = lead;
      else t = mode == 0 ? (lead + lagg) / 2 : (mode == 2 ? (lead < lagg ? lead : lagg) : (lead > lagg ? lead : lagg));
      thr[nt++] = log_mode ? t + scaling : t * scaling;
      /* `if (scaledThr < signal(ix)) peakIndices :+ ix` discards its result (:111): the golden reports no peaks here */
    }
  } else { /* (:115-131) */
    for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
    for (int ix = wc; ix < n - wc; ++ix) {
      const double lead = slice_sum(sig, n, ix - wc, ix - G) / R;
      const double lagg = slice_sum(sig, n, ix + G + 1, ix + wc + 1) / R;
      const double t = mode == 0 ? (lead + lagg) / 2 : (mode == 2 ? (lead < lagg ? lead : lagg) : (lead > lagg ? lead : lagg));
      const double st = log_mode ? t + scaling : t * scaling;
      if (st < sig[ix]) peaks[np++] = ix;
      thr[nt++] = st;
    }
    for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
  }
  *n_thr = nt; *n_peaks = np;
  return 0;
}

/* cfarOS (CFARUtils.scala:159-196): ascending sort of lead ++ lagg, element (1.5*R).toInt; peaks appended twice (:172,174) */
int rspo_golden_cfar_os(const double *sig, int n, int R, int G, double scaling, int log_mode, double *thr, int *n_thr,
                        int *peaks, int *n_peaks) {
  const int wc = R + G;
  int nt = 0, np = 0;
  double *buf = (double *)malloc(sizeof(double) * 2 * (size_t)R);
  if (!buf) return -2;
  for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
  for (int ix = wc; ix < n - wc; ++ix) {
    for (int j = 0; j < R; ++j) { buf[j] = sig[ix - wc + j]; buf[R + j] = sig[ix + G + 1 + j]; }
    qsort(buf, 2 * (size_t)R, sizeof(double), cmp_asc_f64);
    const double t = buf[(int)(1.5 * R)];
    const double st = log_mode ? t + scaling : t * scaling;
    if (st < sig[ix]) peaks[np++] = ix;
    thr[nt++] = st;
    if (st < sig[ix]) peaks[np++] = ix;
  }
  for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
  free(buf);
  *n_thr = nt; *n_peaks = np;
  return 0;
}

/* cfarGOS (CFARUtils.scala:199-272).  Edge padding is appended in BOTH branches (:268-269), as the golden does. */
int rspo_golden_cfar_gos(const double *sig, int n, int R, int G, int index_lagg, int index_lead, int mode,
                         int consider_edges, double scaling, int log_mode, double *thr, int *n_thr, int *peaks,
                         int *n_peaks) {
  const int wc = R + G;
  int nt = 0, np = 0;
  double *lead = (double *)malloc(sizeof(double) * 2 * (size_t)R);
  if (!lead) return -2;
  double *lagg = lead + R;
  for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
  const int lo = consider_edges ? 0 : wc, hi = consider_edges ? n : n - wc;
  for (int ix = lo; ix < hi; ++ix) {
    const int has_lead = !(ix < wc), has_lagg = !(ix >= n - wc);
    if (has_lead) { for (int j = 0; j < R; ++j) lead[j] = sig[ix - wc + j]; qsort(lead, R, sizeof(double), cmp_desc_f64); }
    if (has_lagg) { for (int j = 0; j < R; ++j) lagg[j] = sig[ix + G + 1 + j]; qsort(lagg, R, sizeof(double), cmp_desc_f64); }
    double t;
    if (!has_lead) t = lagg[index_lagg - 1];
    else if (!has_lagg) t = lead[index_lead - 1];
    else {
      const double a = lead[index_lead - 1], b = lagg[index_lagg - 1];
      t = mode == 0 ? (a + b) / 2 : (mode == 2 ? (a < b ? a : b) : (a > b ? a : b));
    }
    const double st = log_mode ? t + scaling : t * scaling;
    if (!consider_edges) { if (st < sig[ix]) peaks[np++] = ix; }
    thr[nt++] = st;
    if (!consider_edges) { if (st < sig[ix]) peaks[np++] = ix; }
  }
  for (int i = 0; i < wc; ++i) thr[nt++] = 0.0;
  free(lead);
  *n_thr = nt; *n_peaks = np;
  return 0;
}

/* cfarCASH (CFARUtils.scala:29-81).  The lagging sub-window end uses `ix - guardCells` (:45), kept as written. */
int rspo_golden_cfar_cash(const double *sig, int n, int R, int sub, int G, double scaling, int log_mode, double *thr,
                          int *n_thr, int *peaks, int *n_peaks) {
  if (!(R > 0 && sub > 0)) return -1;
  const int num_subs = R / sub, wc = R + G;
  int nt = 0, np = 0;
  for (int i = 0; i < R; ++i) thr[nt++] = 0.0; /* edges = Seq.fill(referenceCells) (:62) */
  for (int ix = wc; ix < n - wc; ++ix) {
    double mn = 0.0; int have = 0;
    for (int s = num_subs; s >= 1; --s) {
      const int lead_start = ix - s * sub - G, lead_end = ix - (s - 1) * sub - G;
      const int lagg_start = ix + G + (num_subs - s) * sub + 1, lagg_end = ix - G + (num_subs - s + 1) * sub + 1;
      const double la = slice_sum(sig, n, lead_start, lead_end) / sub;
      const double lg = slice_sum(sig, n, lagg_start, lagg_end) / sub;
      const double mx = la > lg ? la : lg;
      if (!have || mx < mn) { mn = mx; have = 1; }
    }
    const double st = log_mode ? mn + scaling : mn * scaling;
    thr[nt++] = st;
    if (st < sig[ix]) peaks[np++] = ix;
  }
  for (int i = 0; i < R; ++i) thr[nt++] = 0.0;
  *n_thr = nt; *n_peaks = np;
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * PLFG (PLFGDspBlockMem.scala:168-178,227-345; golden plfg_script.py:21-57)
 * ---------------------------------------------------------------------------------------------- */

int64_t rspo_plfg(const rspo_plfg_cfg *c, int32_t *out, int64_t n_out) {
  const int ow = c->output_width, nsw = c->max_num_of_samples_width;
  const int owi = c->output_width_int > 0 ? c->output_width_int : ow;  /* field offsets come from outputWidthInt (:161-171) */
  /* startingPoint is UInt(outputWidthInt.W), slopes is the low outputWidthInt bits of the word: asTypeOf(protoOut) zero-extends both
   * into the (Int + Frac)-bit output and reads the result as a signed raw value (same width when Frac = 0: plain reinterpretation) */
  const int64_t start = rspo_wrap((int64_t)((uint64_t)(uint32_t)c->starting_point & ((1ULL << owi) - 1)), ow);
  /* pause samples between frames: RTL (:242-275) inter - 1, the Python golden (plfg_script.py:50-53) inter */
  const int64_t pause = c->pause_model == 1 ? c->interframe_num_of_samples
                                            : (c->interframe_num_of_samples > 1 ? c->interframe_num_of_samples - 1 : 0);
  int64_t cur = start, k = 0;
  for (int64_t frame = 0; c->frame_num == 0 || frame < c->frame_num; ++frame) {
    for (int d = 0; d < (c->different_chirps_num < 1 ? 1 : c->different_chirps_num); ++d) {
      const uint32_t chirp = c->chirp_ordinal_nums[d];
      const uint32_t reps = c->repeated_chirp_nums[d] < 1 ? 1 : c->repeated_chirp_nums[d]; /* (cnt+1) >= n (:259) */
      for (uint32_t rep = 0; rep < reps; ++rep) {
        const uint32_t nseg = c->segment_nums[chirp] < 1 ? 1 : c->segment_nums[chirp];
        for (uint32_t seg = 0; seg < nseg; ++seg) {
          const uint64_t wd = c->ram[chirp * (uint32_t)c->max_num_of_segments + seg]; /* r_addr (:383) */
          const int64_t slope = rspo_wrap((int64_t)(wd & ((1ULL << owi) - 1)), ow);   /* slopes.asTypeOf(protoOut) */
          uint32_t ns = (uint32_t)((wd >> owi) & ((1ULL << nsw) - 1));
          if (!(ns > 1)) ns += 1;                                                      /* (:175) */
          const int sign = (int)((wd >> (owi + nsw)) & 1), reset = (int)((wd >> (owi + nsw + 1)) & 1);
          for (uint32_t s = 0; s < ns; ++s) {
            if (reset || (seg == 0 && s == 0)) cur = start;                            /* plfg_script.py:46-47 */
            else cur = rspo_wrap(sign ? cur + slope : cur - slope, ow);                /* (:343) */
            if (k >= n_out) return k;
            out[k++] = (int32_t)cur;
          }
        }
      }
    }
    if (c->frame_num == 0 || frame < c->frame_num - 1) { /* the inter-frame pause emits startingPoint */
      for (int64_t i = 0; i < pause; ++i) {
        cur = start;
        if (k >= n_out) return k;
        out[k++] = (int32_t)cur;
      }
    }
  }
  return k;
}

/* ------------------------------------------------------------------------------------------------
 * NCO (NCOLazyModuleBlock.scala:313-324,450-455; NCOTable.scala:105-113,193-280)
 * ---------------------------------------------------------------------------------------------- */

uint32_t rspo_nco_words(const rspo_nco_cfg *c, const int32_t *freq, uint32_t *out, int64_t n) {
  const int pw = c->phase_width, ab = ilog2_ceil((uint32_t)c->table_size), ts = c->table_size;
  const uint32_t pmask = pw >= 32 ? 0xFFFFFFFFu : ((1u << pw) - 1);
  const int32_t one = 1 << (c->table_width - 2);
  int32_t *tab = (int32_t *)malloc(sizeof(int32_t) * (size_t)ts);
  if (!tab) return 0;
  rspo_nco_table(ts, c->table_width, tab);
  uint32_t phase = c->phase0 & pmask;
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t f = (uint32_t)freq[i] & pmask;
    phase = c->phase_acc_enable ? ((phase + f) & pmask) : f; /* (:313-321) */
    const uint32_t ph = (phase + c->poff) & pmask;           /* (:324) */
    uint32_t msbs, addr, raddr;
    if (c->rasterized) {
      /* NCOTableRasterizedMode (:334-336): addrR = phase % tableSize, msbs = (phase / tableSize) % 4; addrReverse = tableSize - addr (:356-358),
       * which for addr 0 points past the table: that look-up is replaced by 1.0 (cosIsOne / sinIsOne, :385-413) */
      addr = ph % (uint32_t)ts;
      msbs = (ph / (uint32_t)ts) % 4u;
      raddr = (uint32_t)ts - addr;
      if (raddr >= (uint32_t)ts) raddr = 0; /* only with addr == 0, where the value is not used */
    } else {
      msbs = ph >> (pw - 2);
      addr = (ph >> (pw - 2 - ab)) & (uint32_t)(ts - 1);
      raddr = (0u - addr) & (uint32_t)(ts - 1); /* addrReverse (:172-174) */
    }
    int32_t s, co;
    switch (msbs) { /* (:193-235), reversed lookup replaced by 1.0 when addr == 0 (:201,211,262-263) */
    case 0: s = tab[addr]; co = (addr == 0) ? one : tab[raddr]; break;
    case 1: s = (addr == 0) ? one : tab[raddr]; co = -tab[addr]; break;
    case 2: s = -tab[addr]; co = -((addr == 0) ? one : tab[raddr]); break;
    default: s = -((addr == 0) ? one : tab[raddr]); co = tab[addr]; break;
    }
    if (c->use_multiplier) { /* (NCOLazyModuleBlock.scala:354-364) beatBytes = 4 */
      const int w = c->table_width;
      const int64_t f = rspo_wrap(c->enable_multiplying ? (int64_t)c->multiplying_factor : ((int64_t)1 << 14), w);
      s = (int32_t)rspo_wrap(rspo_trim((int64_t)s * f, w - 2, c->rounding_mode), w);
      co = (int32_t)rspo_wrap(rspo_trim((int64_t)co * f, w - 2, c->rounding_mode), w);
    }
    out[i] = (((uint32_t)co & 0xFFFFu) << 16) | ((uint32_t)s & 0xFFFFu); /* Cat(cos, sin) (:452) */
  }
  free(tab);
  return phase;
}

/* ------------------------------------------------------------------------------------------------
 * whole chain (FftMagCfarChain.scala:47): fft -> mag -> cfar on 32-bit beats
 * ---------------------------------------------------------------------------------------------- */

int rspo_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

int rspo_chain_fixed_words(const rspo_chain_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps, int n_threads,
                           uint32_t *fft_out, uint32_t *mag_out) {
  const int N = 1 << cfg->fft.log_n;
  if (cfg->cfar.fft_win != N) return -1;
  int rc_all = 0;
  (void)n_threads;
  if (cfg->fft.data_width > 16 || cfg->fft.log_n < 1 || cfg->fft.log_n > 24) return -1;
  fft_tabs tabs;
  if (fft_tabs_build(&tabs, &cfg->fft)) return -2;
#ifdef _OPENMP
#pragma omp parallel num_threads(n_threads > 0 ? n_threads : 1)
#endif
  {
    uint32_t *f = (uint32_t *)malloc(sizeof(uint32_t) * 2 * (size_t)N);
    uint32_t *m = f ? f + N : NULL;
    int rc = f ? 0 : -2;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
    for (int c = 0; c < n_chirps; ++c) {
      if (rc) continue;
      const uint32_t *src = in + (size_t)c * N;
      rc = fft_fixed_words_tabs(&cfg->fft, &tabs, src, f, 1);
      if (!rc) rc = rspo_mag_words(&cfg->mag, f, m, N);
      if (!rc) rc = rspo_cfar_words(&cfg->cfar, m, out + (size_t)c * N, 1);
      if (!rc && fft_out) memcpy(fft_out + (size_t)c * N, f, sizeof(uint32_t) * N);
      if (!rc && mag_out) memcpy(mag_out + (size_t)c * N, m, sizeof(uint32_t) * N);
    }
    if (rc) {
#ifdef _OPENMP
#pragma omp critical
#endif
      rc_all = rc;
    }
    free(f);
  }
  fft_tabs_free(&tabs);
  return rc_all;
}

/* ------------------------------------------------------------------------------------------------
 * java.util.Random
 * ---------------------------------------------------------------------------------------------- */
void rspo_jrand_seed(rspo_jrand *r, int64_t seed) { r->seed = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
static int32_t jrand_next(rspo_jrand *r, int bits) {
  r->seed = (r->seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
  return (int32_t)((int64_t)r->seed >> (48 - bits));
}
double rspo_jrand_next_double(rspo_jrand *r) {
  const int64_t a = jrand_next(r, 26), b = jrand_next(r, 27);
  return (double)((a << 27) + b) * (1.0 / 9007199254740992.0);
}
