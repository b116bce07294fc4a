// rsp_launch.hpp -- host-side declarations of the kernel launchers.  The templated kernels are instantiated in several
// translation units (rsp_inst.cu compiled once per RSP_UNIT, rsp_fast.cu) so the library builds in parallel.
#pragma once
#include <cuda_runtime.h>

#include "../../include/rsp_chains.h"
#include "rsp_device.cuh"
#include "rsp_device_f32.cuh"

namespace rsp {

struct LaunchCtx { int device, sm_count; };

// RSP_OK, or a status with *err pointing at a static message
#define RSP_DECL_FIXED(unit) rsp_status launch_fixed_unit##unit(int mode, int S, const LaunchCtx &, const FxArgs &, cudaStream_t, const char **err)
RSP_DECL_FIXED(0); RSP_DECL_FIXED(1); RSP_DECL_FIXED(2); RSP_DECL_FIXED(3); RSP_DECL_FIXED(4); RSP_DECL_FIXED(5); RSP_DECL_FIXED(8); RSP_DECL_FIXED(9);
RSP_DECL_FIXED(10); RSP_DECL_FIXED(11); RSP_DECL_FIXED(12); RSP_DECL_FIXED(13);
#undef RSP_DECL_FIXED
rsp_status launch_f32_unit6(int S, const LaunchCtx &, const F32Args &, cudaStream_t, const char **err);
rsp_status launch_f32_unit7(int S, const LaunchCtx &, const F32Args &, cudaStream_t, const char **err);

// units: fixed kernels 0..5, 8..13 = (mode, S range); fp32 kernels 6, 7 = S range
constexpr int kSplitS = 10;   // S < kSplitS -> low unit, else high unit
inline rsp_status launch_fixed_inst(int mode, int S, const LaunchCtx &c, const FxArgs &a, cudaStream_t st, const char **err) {
  const bool hi = S >= kSplitS;
  switch (mode) {
  case MODE_FAST: return hi ? launch_fixed_unit3(mode, S, c, a, st, err) : launch_fixed_unit2(mode, S, c, a, st, err);
  case MODE_FAST3: return hi ? launch_fixed_unit5(mode, S, c, a, st, err) : launch_fixed_unit4(mode, S, c, a, st, err);
  case MODE_RADIX2: return hi ? launch_fixed_unit9(mode, S, c, a, st, err) : launch_fixed_unit8(mode, S, c, a, st, err);
  case MODE_FASTC: return hi ? launch_fixed_unit11(mode, S, c, a, st, err) : launch_fixed_unit10(mode, S, c, a, st, err);
  case MODE_FAST3C: return hi ? launch_fixed_unit13(mode, S, c, a, st, err) : launch_fixed_unit12(mode, S, c, a, st, err);
  default: return hi ? launch_fixed_unit1(mode, S, c, a, st, err) : launch_fixed_unit0(mode, S, c, a, st, err);
  }
}
inline rsp_status launch_f32_inst(int S, const LaunchCtx &c, const F32Args &a, cudaStream_t st, const char **err) {
  return S >= kSplitS ? launch_f32_unit7(S, c, a, st, err) : launch_f32_unit6(S, c, a, st, err);
}

template <typename K>
inline rsp_status launch_cfg(K kernel, size_t smem, int threads, int *blocks_per_sm, const char **err) {
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    *err = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed"; return RSP_ECUDA; }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem) != cudaSuccess || occ < 1) {
    *err = "kernel does not fit on an SM"; return RSP_ECUDA; }
  *blocks_per_sm = occ;
  return RSP_OK;
}

}  // namespace rsp
