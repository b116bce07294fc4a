// rsp_kernels.cuh -- the CUDA kernels (sm_100a) built from the per-thread phases of rsp_device.cuh / rsp_device_f32.cuh.
//
// Layout.  A "row" is one chirp of one RX channel: N = 2^S beats, rows back to back in HBM.  A CTA owns CPC rows at a
// time (TPC threads per row, 16 points per thread) and walks the grid-stride list of row groups, so a grid of
// (#SMs x resident CTAs) stays busy for any batch size.  A row lives in shared memory from its first global load to its
// last global store: FFT passes ping through bufA, the natural-order magnitude beats land in bufM (the bit reversal is the
// scatter into bufM), the CFAR prefix sums reuse bufA.  Global loads and stores are warp-contiguous.
#pragma once
#include <cuda_runtime.h>

#include "rsp_device.cuh"
#include "rsp_device_f32.cuh"
#include "rsp_device_gen.cuh"

namespace rsp {

// exclusive scan of one int per thread across the TPC threads of a row (warp shuffles + one smem hop across warps)
template <int S>
__device__ __forceinline__ int row_scan_exclusive(int total, int c, int tl, int *wsum) {
  constexpr int GRP = Geo<S>::GRP, NGRP = Geo<S>::NGRP;
  const int lane = tl & (GRP - 1);
  int incl = total;
#pragma unroll
  for (int d = 1; d < GRP; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d, GRP);
    if (lane >= d) incl += v;
  }
  int offset = 0;
  if (NGRP > 1) {
    const int g = tl / GRP;
    if (lane == GRP - 1) wsum[c * NGRP + g] = incl;
    __syncthreads();
    for (int h = 0; h < g; ++h) offset += wsum[c * NGRP + h];
  }
  return offset + incl - total;
}

template <int S, int MODE, int K>
__device__ __forceinline__ void run_fft_passes(const FxArgs &a, bool active, int tl, const uint32_t *gin, uint32_t *bufA, uint32_t *bufM) {
  if (active) fft_pass<S, MODE, K>(a, tl, gin, bufA, bufM);
  __syncthreads();
  if constexpr (K + 1 < fx_n_passes(S)) run_fft_passes<S, MODE, K + 1>(a, active, tl, gin, bufA, bufM);
}

// ---------------------------------------------------------------------------------------------------------------------
// fixed point: [FFT] -> [LogMagMux] -> [CFAR], stages selected by a.first_stage / a.last_stage
template <int S, int MODE>
__global__ void __launch_bounds__(Geo<S>::THREADS) chain_fixed_kernel(const FxArgs a) {
  using G = Geo<S>;
  constexpr int N = G::N, TPC = G::TPC, CPC = G::CPC, PTS = G::PTS, BUF = G::BUF;
  extern __shared__ uint32_t smem[];
  __shared__ int wsum[CPC * G::NGRP];
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  uint32_t *bufA = smem + (size_t)c * 2 * BUF, *bufM = bufA + BUF, *bufP = bufA;
  for (uint32_t row0 = blockIdx.x * CPC; row0 < a.n_rows; row0 += gridDim.x * CPC) {
    const uint32_t row = row0 + c;
    const bool active = row < a.n_rows;
    const uint32_t *gin = a.in + (size_t)row * N;
    if (a.first_stage == STAGE_FFT) {
      run_fft_passes<S, MODE, 0>(a, active, tl, gin, bufA, bufM);
    } else {
      if (active) {
#pragma unroll
        for (int j = 0; j < PTS; ++j) { const int n = tl + j * TPC; bufM[pad(n)] = ldg(&gin[n]); }
      }
      __syncthreads();
    }
    if (a.last_stage != STAGE_CFAR) {
      if (active) {
        uint32_t *gout = a.out + (size_t)row * N;
#pragma unroll
        for (int j = 0; j < PTS; ++j) { const int n = tl + j * TPC; gout[n] = bufM[pad(n)]; }
      }
      __syncthreads();
      continue;
    }
    const bool need_sums = !((a.core == CORE_LIS && a.cfar_algorithm == 1) || a.core == CORE_OS);
    if (need_sums) {
      const int total = cfar_thread_total<S>(a, tl, bufM);
      const int offset = row_scan_exclusive<S>(total, c, tl, wsum);
      cfar_write_prefix<S>(a, tl, offset, bufM, bufP);
      __syncthreads();
    }
    if (active) {
#pragma unroll 4
      for (int j = 0; j < PTS; ++j) cfar_bin<S>(a, row, tl + j * TPC, bufM, bufP, a.out);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// fp32 flavour
template <int S, int K>
__device__ __forceinline__ void run_fft_passes_f32(const F32Args &a, bool active, int tl, const F2 *gin, F2 *bufA, float *bufM, F2 *gout_c) {
  if (active) fft_pass_f32<S, K>(a, tl, gin, bufA, bufM, gout_c);
  __syncthreads();
  if constexpr (K + 1 < fx_n_passes(S)) run_fft_passes_f32<S, K + 1>(a, active, tl, gin, bufA, bufM, gout_c);
}

template <int S>
__global__ void __launch_bounds__(Geo<S>::THREADS) chain_f32_kernel(const F32Args a) {
  using G = GeoF<S>;
  constexpr int N = G::N, TPC = G::TPC, PTS = G::PTS, BUFA = G::BUFA, BUFM = G::BUFM, MASKW = G::MASKW;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int c = threadIdx.x / TPC, tl = threadIdx.x % TPC;
  // per row: bufA (BUFA complex) then bufM (BUFM floats, rounded up to keep bufA 8-byte aligned); thresholds reuse bufA
  constexpr size_t ROW_BYTES = (size_t)BUFA * 8 + (((size_t)BUFM * 4 + 7) & ~(size_t)7);
  F2 *bufA = reinterpret_cast<F2 *>(smem_raw + (size_t)c * ROW_BYTES);
  float *bufM = reinterpret_cast<float *>(smem_raw + (size_t)c * ROW_BYTES + (size_t)BUFA * 8);
  float *bufT = reinterpret_cast<float *>(bufA);
  uint32_t *bufB = reinterpret_cast<uint32_t *>(bufT + BUFM);   // the threads' peak bits, TPC words (fits: BUFA*8 >= BUFM*4 + TPC*4)
  for (uint32_t row0 = blockIdx.x * G::CPC; row0 < a.n_rows; row0 += gridDim.x * G::CPC) {
    const uint32_t row = row0 + c;
    const bool active = row < a.n_rows;
    if (a.first_stage == STAGE_FFT) {
      run_fft_passes_f32<S, 0>(a, active, tl, a.in + (size_t)row * N, bufA, bufM, a.out_c ? a.out_c + (size_t)row * N : nullptr);
      if (a.last_stage == STAGE_FFT) continue;
    } else {
      if (active) {
        const float *gin = a.in_mag + (size_t)row * N;
#pragma unroll
        for (int j = 0; j < PTS; ++j) { const int n = tl + j * TPC; bufM[pad(n)] = gin[n]; }
      }
      __syncthreads();
    }
    if (a.last_stage != STAGE_CFAR) {
      if (active) {
        float *gout = a.out + (size_t)row * N;
#pragma unroll
        for (int j = 0; j < PTS; ++j) { const int n = tl + j * TPC; gout[n] = bufM[pad(n)]; }
      }
      __syncthreads();
      continue;
    }
    uint32_t bits = 0;
    if (active) bits = cfar_thread_f32<S>(a, row, tl, bufM, bufT);
    bufB[tl] = bits;
    __syncthreads();
    if (active) {
      if (a.out) {
        float *gout = a.out + (size_t)row * N;
#pragma unroll
        for (int j = 0; j < PTS; ++j) { const int n = tl + j * TPC; gout[n] = bufT[pad(n)]; }
      }
      if (a.peak_mask) {
        // word w of the row's mask covers bins 32w..32w+31 = the PTS-bit fields of 32/PTS consecutive threads
        constexpr int TPW = 32 / PTS > 0 ? 32 / PTS : 1;   // threads per mask word (PTS <= 16 -> >= 2)
        for (int w = tl; w < MASKW; w += TPC) {
          uint32_t m = 0;
#pragma unroll
          for (int q = 0; q < TPW; ++q) {
            const int t = w * TPW + q;
            if (t < TPC) m |= bufB[t] << (q * PTS);
          }
          if (N < 32) m &= (1u << (N & 31)) - 1u;
          a.peak_mask[(size_t)row * MASKW + w] = m;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace rsp
