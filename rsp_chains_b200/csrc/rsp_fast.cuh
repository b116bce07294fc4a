// rsp_fast.cuh -- specialised kernels for the chain configurations the benchmarks run (selected by fast_chain_supported).
#pragma once
#include <cuda_runtime.h>

#include "../../include/rsp_chains.h"
#include "rsp_host.hpp"
#include "rsp_launch.hpp"

// fixed-point CA chain and FFT block: 64 ... 16384 points (the OS / GOS kernels and the fp32 flavour from 1024 on)
#ifndef RSP_FAST_SMIN
#define RSP_FAST_SMIN 6
#endif
#ifndef RSP_FAST_SMAX
#define RSP_FAST_SMAX 14
#endif

namespace rsp {
bool fast_chain_supported(const Plan &pl, const FxArgs &a);
rsp_status launch_fast_chain(const Plan &pl, const LaunchCtx &ctx, const FxArgs &a, cudaStream_t st, const char **err);
// the FFT block alone (rsp_fft_run) on the fast passes
bool fast_fft_supported(const Plan &pl, const FxArgs &a);
rsp_status launch_fast_fft(const Plan &pl, const LaunchCtx &ctx, const FxArgs &a, cudaStream_t st, const char **err);
// fp32 flavour (rsp_fast_f32.cu); tw4 = Tables::twf4 on the device
bool fast_f32_chain_supported(const Plan &pl, const F32Args &a);
rsp_status launch_fast_f32_chain(const Plan &pl, const LaunchCtx &ctx, const F32Args &a, const float4 *tw4, cudaStream_t st, const char **err);
}  // namespace rsp
