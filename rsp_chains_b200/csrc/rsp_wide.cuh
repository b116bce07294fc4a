// rsp_wide.cuh -- the FFT block with wide beats (rsp_wide.cu): arguments and launcher
#pragma once
#include <cuda_runtime.h>

#include "../../include/rsp_chains.h"
#include "rsp_device.cuh"
#include "rsp_launch.hpp"

namespace rsp {

struct WideArgs {
  const unsigned long long *in;   // 64-bit beats, low 2 * w0 bits = Cat(re, im)
  unsigned long long *out;        // out_words uint64 per beat: Cat(re, im) with wo-bit halves, low word first
  uint32_t n_rows;
  int S, w0, wo, out_words;       // transform size, dataWidth, width of the last stage's type
  int bp, tw_bp, trim, use4muls, bitrev, fft_dir;
  uint32_t keep_mask, expand_mask;   // per stage of THIS transform
  int base_growth;                // bits added by the physical stages in front of this transform (run-time size)
  const int2 *tw;                 // raw (cos, -sin) as int32 pairs in Q2.(twiddleWidth - 2), per stageN three runs [W^2k | W^k | W^3k]
  int tw_off[16];
  const int32_t *win;
  int win_on, win_shift;
  uint32_t *ovf_flags;
  int ovf_shift;
};

rsp_status launch_fft_wide(const LaunchCtx &ctx, const WideArgs &a, cudaStream_t st, const char **err);

}  // namespace rsp
