// rsp_wide.cu -- the FFT block with beats wider than Cat(re16, im16): data or twiddles wider than 16 bits and / or grow stages
// (`expandLogic`), the configurations behind the reference's 24-bit and "grow" SQNR results
// (generators/sdf-fft/src/test/scala/Radix22SQNRSpec.scala:60-150).  Radix-2^2 DIF chain (SDFChainRadix22.scala:224-352,453-493) in
// 64-bit integer arithmetic: stage i works in the type protoIQstages(i) = FixedPoint(dataWidth + sum expandLogic(0..i), binPoint)
// (SDFFFTParams.scala:119-123); a grow stage keeps the grown butterfly outputs, the others halve them under trimType or keep the LSBs.
// A parity path, not a fast one: one CTA per row, the row in shared memory as 64-bit (re, im) pairs, one barrier per stage.
#include <algorithm>
#include <cstdint>

#include "rsp_wide.cuh"

namespace rsp {

__device__ __forceinline__ void wide_cmul(const WideArgs &a, int w, long long &xr, long long &xi, int2 tw) {
  // DspComplex context_* (trimType, Grow) then trimBinary(bpos), assigned to the stage's type (SDFChainRadix22.scala:298-308)
  const long long br = tw.x, bi = tw.y, ar = xr, ai = xi;
  const int bp_a = a.bp, bp_b = a.tw_bp, trim = a.trim;
  const int prod_bp = bp_a + bp_b;
  const int ctx_bp = (bp_a > bp_b ? bp_a : bp_b) + 1;
  const int s1 = prod_bp > ctx_bp ? prod_bp - ctx_bp : 0;
  const int cur_bp = prod_bp - s1;
  long long re, im;
  if (a.use4muls) {
    re = trim_rt<long long>(ar * br, s1, trim) - trim_rt<long long>(ai * bi, s1, trim);
    im = trim_rt<long long>(ar * bi, s1, trim) + trim_rt<long long>(ai * br, s1, trim);
  } else {
    const long long k1 = trim_rt<long long>(ar * (br + bi), s1, trim);
    const long long k2 = trim_rt<long long>((ar + ai) * bi, s1, trim);
    const long long k3 = trim_rt<long long>((ai - ar) * br, s1, trim);
    re = k1 - k2;
    im = k1 + k3;
  }
  if (cur_bp > bp_a) { re = trim_rt<long long>(re, cur_bp - bp_a, trim); im = trim_rt<long long>(im, cur_bp - bp_a, trim); }
  else if (cur_bp < bp_a) { re <<= (bp_a - cur_bp); im <<= (bp_a - cur_bp); }
  xr = wrapw64(re, w);
  xi = wrapw64(im, w);
}

// what follows the butterfly of stage i for the point at position p (SDFChainRadix22.scala:224-352)
__device__ __forceinline__ void wide_post(const WideArgs &a, int i, int D, int w, int p, long long &re, long long &im) {
  if (i == a.S - 1) return;
  if ((i & 1) == 0) {   // x(-j) on block positions [3D/2, 2D)
    if ((p & (2 * D - 1)) >= 3 * D / 2) { const long long r = re; re = im; im = wrapw64(-r, w); }
  } else {              // x twiddle [0]*D ++ [2k] ++ [k] ++ [3k] over p mod 4D
    const int pos = p & (4 * D - 1), quarter = pos / D, k = pos & (D - 1);
    if (quarter) {
      const int slot = quarter == 1 ? 0 : (quarter == 2 ? 1 : 2);
      const int L = a.S - i + 1;   // stageN = 4 D = 1 << L
      wide_cmul(a, w, re, im, __ldg(&a.tw[a.tw_off[L] + slot * D + k]));
    }
  }
}

__global__ void __launch_bounds__(256) fft_wide_kernel(const WideArgs a) {
  extern __shared__ __align__(16) long long wsm[];
  const int N = 1 << a.S;
  long long *xr = wsm, *xi = wsm + N;
  const unsigned long long m0 = a.w0 >= 64 ? ~0ull : ((1ull << a.w0) - 1ull);
  for (uint32_t row = blockIdx.x; row < a.n_rows; row += gridDim.x) {
    const unsigned long long *gin = a.in + (size_t)row * N;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
      const unsigned long long wd = gin[p];
      long long re = wrapw64((long long)((wd >> a.w0) & m0), a.w0), im = wrapw64((long long)(wd & m0), a.w0);
      if (a.win_on) {   // in * window(cntWin), truncated on assignment to protoIQ (SDFFFT.scala:233-234)
        const long long wv = __ldg(&a.win[p]);
        re = wrapw64((re * wv) >> a.win_shift, a.w0); im = wrapw64((im * wv) >> a.win_shift, a.w0);
      }
      xr[p] = a.fft_dir ? re : im;   // ifft swaps real and imaginary on the way in (SDFChainRadix22.scala:205-211)
      xi[p] = a.fft_dir ? im : re;
    }
    __syncthreads();
    uint32_t ovf = 0;
    int w = a.w0 + a.base_growth;   // width of the type the stage's inputs arrive in
    for (int i = 0; i < a.S; ++i) {
      const int LD = a.S - 1 - i, D = 1 << LD;
      const bool grow = (a.expand_mask >> i) & 1u, scale = (a.keep_mask >> i) & 1u;
      const int w_in = w;
      w += grow ? 1 : 0;
      const long long lim = 1ll << (w_in - 1);
      for (int k = threadIdx.x; k < N / 2; k += blockDim.x) {
        const int p0 = ((k >> LD) << (LD + 1)) | (k & (D - 1)), p1 = p0 + D;
        const long long ar = xr[p0], ai = xi[p0], cr = xr[p1], ci = xi[p1];
        const long long sr = ar + cr, si = ai + ci, dr = ar - cr, di = ai - ci;
        long long o0r, o0i, o1r, o1i;
        if (grow) { o0r = sr; o0i = si; o1r = dr; o1i = di; }   // kept as they are (SDFChainRadix22.scala:454-457)
        else {
          if (sr >= lim || si >= lim || dr >= lim || di >= lim) ovf |= 1u << i;   // `tooBig` (:473-481)
          if (scale) {
            o0r = wrapw64(trim_rt<long long>(sr, 1, a.trim), w); o0i = wrapw64(trim_rt<long long>(si, 1, a.trim), w);
            o1r = wrapw64(trim_rt<long long>(dr, 1, a.trim), w); o1i = wrapw64(trim_rt<long long>(di, 1, a.trim), w);
          } else {
            o0r = wrapw64(sr, w); o0i = wrapw64(si, w); o1r = wrapw64(dr, w); o1i = wrapw64(di, w);
          }
        }
        wide_post(a, i, D, w, p0, o0r, o0i);
        wide_post(a, i, D, w, p1, o1r, o1i);
        xr[p0] = o0r; xi[p0] = o0i; xr[p1] = o1r; xi[p1] = o1i;
      }
      __syncthreads();
    }
    if (ovf != 0 && a.ovf_flags != nullptr) atomicOr(a.ovf_flags, ovf << a.ovf_shift);
    // out = Cat(re, im) of the last stage's type, natural order through the bit reversal when useBitReverse
    const unsigned long long mo = (1ull << a.wo) - 1ull;
    unsigned long long *gout = a.out + (size_t)row * N * a.out_words;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
      const int p = a.bitrev ? (int)(__brev((uint32_t)n) >> (32 - a.S)) : n;
      const long long vr = a.fft_dir ? xr[p] : xi[p], vi = a.fft_dir ? xi[p] : xr[p];   // ifft swaps again on the way out; scalar 1.0
      const unsigned long long re = (unsigned long long)vr & mo, im = (unsigned long long)vi & mo;
      if (a.out_words == 1) gout[n] = (re << a.wo) | im;
      else { gout[2 * n] = (a.wo < 64 ? (re << a.wo) : 0ull) | im; gout[2 * n + 1] = re >> (64 - a.wo); }
    }
    __syncthreads();
  }
}

rsp_status launch_fft_wide(const LaunchCtx &ctx, const WideArgs &a, cudaStream_t st, const char **err) {
  static int attr_dev[64] = {0};
  const size_t smem = (size_t)2 * sizeof(long long) << a.S;
  if (smem > 48 * 1024 && !attr_dev[ctx.device & 63]) {
    if (cudaFuncSetAttribute(fft_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * (int)sizeof(long long) * 8192) != cudaSuccess) {
      *err = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed"; return RSP_ECUDA; }
    attr_dev[ctx.device & 63] = 1;
  }
  const unsigned grid = std::max(1u, std::min(a.n_rows, (unsigned)(ctx.sm_count * 2)));
  fft_wide_kernel<<<grid, 256, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) { *err = "fft_wide_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}

}  // namespace rsp
