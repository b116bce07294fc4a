// rsp_doppler.cuh -- the two data movements of the slow-time (Doppler) dimension, SURVEY 8(f)2.  Not in the reference (its chains stop at
// the 1-D range CFAR); defined here so that the reference's own blocks -- the same SDF-FFT, LogMagMux and CFAR, instantiated a second time
// with numPoints = chirps per frame -- can run along slow time:
//   corner_turn_kernel   out[b][c][r] = in[b][r][c]                     (range-major -> Doppler-major and back), HBM bound, 2 x elem bytes per cell
//   mag_combine_*_kernel out[i] = mean over the RX channels of LogMagMux(in[rx][i])   (non-coherent combining), (4 n_rx + 4) B per cell
#pragma once
#include <cstdint>

#include "rsp_device.cuh"
#include "rsp_device_f32.cuh"

namespace rsp {

// One TILE x TILE tile per CTA iteration through shared memory, reads and writes both warp-contiguous.  256 threads; tile rows are padded by
// PAD elements so that the column reads are conflict free (one word for 2- and 4-byte elements, one element for 8-byte ones).
template <typename T, int TILE>
__global__ void __launch_bounds__(256) corner_turn_kernel(const T *__restrict__ in, T *__restrict__ out, uint32_t n_rows, uint32_t n_cols,
                                                          uint32_t tiles_r, uint32_t tiles_c, unsigned long long n_tiles) {
  constexpr int PAD = sizeof(T) == 2 ? 2 : 1, RPP = 256 / TILE;   // tile rows covered by one pass of the CTA
  __shared__ T tile[TILE][TILE + PAD];
  const int tx = threadIdx.x % TILE, ty = threadIdx.x / TILE;
  const unsigned long long per_batch = (unsigned long long)tiles_r * tiles_c;
  for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const unsigned long long b = t / per_batch, w = t % per_batch;
    const uint32_t r0 = (uint32_t)(w / tiles_c) * TILE, c0 = (uint32_t)(w % tiles_c) * TILE;
    const T *src = in + b * n_rows * (unsigned long long)n_cols;
    T *dst = out + b * n_rows * (unsigned long long)n_cols;
#pragma unroll
    for (int j = 0; j < TILE; j += RPP) {
      const uint32_t r = r0 + ty + j, c = c0 + tx;
      if (r < n_rows && c < n_cols) tile[ty + j][tx] = __ldg(&src[(unsigned long long)r * n_cols + c]);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < TILE; j += RPP) {
      const uint32_t c = c0 + ty + j, r = r0 + tx;
      if (r < n_rows && c < n_cols) dst[(unsigned long long)c * n_rows + r] = tile[tx][ty + j];
    }
    __syncthreads();
  }
}

// LogMagMux (the handle's magnitude registers: sel 1 / 2 / 3) of every RX channel's beat i, summed and divided by n_rx = 2^shift (truncating
// shift, as the CFAR cores divide their window sums, CFARCoreWithMem.scala:132-143); out beat = Cat(0, mean) like LogMagMux's own.
__global__ void __launch_bounds__(256) mag_combine_fixed_kernel(const FxArgs a, unsigned long long n_beats, uint32_t n_rx, int shift) {
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_beats; i += stride) {
    long long sum = 0;
    for (uint32_t rx = 0; rx < n_rx; ++rx) {
      const uint32_t wd = ldg(&a.in[rx * n_beats + i]);
      const int re = wrapw((int)(wd >> 16), a.mag_w), im = wrapw((int)(wd & 0xFFFFu), a.mag_w);
      int ore;
      sum += mag_one(a, re, im, &ore);
    }
    a.out[i] = pack_iq(0, (int)(sum >> shift));
  }
}

__global__ void __launch_bounds__(256) mag_combine_f32_kernel(const F32Args a, unsigned long long n_beats, uint32_t n_rx, float inv) {
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_beats; i += stride) {
    float sum = 0.f;
    for (uint32_t rx = 0; rx < n_rx; ++rx) { const F2 v = a.in[rx * n_beats + i]; sum = __fadd_rn(sum, mag_f32(a.sel, v.x, v.y)); }
    a.out[i] = __fmul_rn(sum, inv);
  }
}

}  // namespace rsp
