// rsp_inst.cu -- instantiates one slice of the templated chain kernels.  Compiled once per RSP_UNIT (see Makefile):
//   0/1 generic fixed, 2/3 fast fixed, 4/5 fast3 fixed, 8/9 radix-2 fixed, 10/11 fast Convergent, 12/13 fast3 Convergent (low / high transform sizes);
//   6/7 fp32 (low / high).
#include <algorithm>

#include "rsp_kernels.cuh"
#include "rsp_launch.hpp"

#ifndef RSP_UNIT
#error "compile with -DRSP_UNIT=0..13"
#endif

namespace rsp {

#if RSP_UNIT < 6 || RSP_UNIT >= 8
template <int S, int MODE>
static rsp_status launch_fixed_sm(const LaunchCtx &c, const FxArgs &a, cudaStream_t st, const char **err) {
  using G = Geo<S>;
  static int occ_dev[64] = {0};   // per instantiation and device: the shared-memory attribute is per context, set once
  int &occ = occ_dev[c.device & 63];
  const size_t smem = (size_t)G::CPC * 2 * G::BUF * sizeof(uint32_t);
  auto kernel = chain_fixed_kernel<S, MODE>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, G::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (a.n_rows + G::CPC - 1) / G::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, G::THREADS, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_fixed_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}
#define RSP_CASE(s) case s: return launch_fixed_sm<s, RSP_MODE_OF_UNIT>(c, a, st, err);
#if RSP_UNIT / 2 == 0
#define RSP_MODE_OF_UNIT MODE_GENERIC
#elif RSP_UNIT / 2 == 1
#define RSP_MODE_OF_UNIT MODE_FAST
#elif RSP_UNIT / 2 == 4
#define RSP_MODE_OF_UNIT MODE_RADIX2
#elif RSP_UNIT / 2 == 5
#define RSP_MODE_OF_UNIT MODE_FASTC
#elif RSP_UNIT / 2 == 6
#define RSP_MODE_OF_UNIT MODE_FAST3C
#else
#define RSP_MODE_OF_UNIT MODE_FAST3
#endif
#define RSP_CAT2(a, b) a##b
#define RSP_CAT(a, b) RSP_CAT2(a, b)
rsp_status RSP_CAT(launch_fixed_unit, RSP_UNIT)(int, int S, const LaunchCtx &c, const FxArgs &a, cudaStream_t st, const char **err) {
  switch (S) {
#if RSP_UNIT % 2 == 0
    RSP_CASE(1) RSP_CASE(2) RSP_CASE(3) RSP_CASE(4) RSP_CASE(5) RSP_CASE(6) RSP_CASE(7) RSP_CASE(8) RSP_CASE(9)
#else
    RSP_CASE(10) RSP_CASE(11) RSP_CASE(12) RSP_CASE(13) RSP_CASE(14)
#endif
  default: *err = "transform size must be 2..16384"; return RSP_EUNSUPPORTED;
  }
}
#else
template <int S>
static rsp_status launch_f32_s(const LaunchCtx &c, const F32Args &a, cudaStream_t st, const char **err) {
  using G = GeoF<S>;
  static int occ_dev[64] = {0};
  int &occ = occ_dev[c.device & 63];
  const size_t row_bytes = (size_t)G::BUFA * 8 + (((size_t)G::BUFM * 4 + 7) & ~(size_t)7);
  const size_t smem = (size_t)G::CPC * row_bytes;
  auto kernel = chain_f32_kernel<S>;
  if (!occ) { const rsp_status s = launch_cfg(kernel, smem, G::THREADS, &occ, err); if (s != RSP_OK) return s; }
  const unsigned groups = (a.n_rows + G::CPC - 1) / G::CPC;
  const unsigned grid = std::max(1u, std::min(groups, (unsigned)(c.sm_count * occ)));
  kernel<<<grid, G::THREADS, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) { *err = "chain_f32_kernel launch failed"; return RSP_ECUDA; }
  return RSP_OK;
}
#define RSP_CASE(s) case s: return launch_f32_s<s>(c, a, st, err);
#if RSP_UNIT == 6
rsp_status launch_f32_unit6(int S, const LaunchCtx &c, const F32Args &a, cudaStream_t st, const char **err) {
  switch (S) {
    RSP_CASE(1) RSP_CASE(2) RSP_CASE(3) RSP_CASE(4) RSP_CASE(5) RSP_CASE(6) RSP_CASE(7) RSP_CASE(8) RSP_CASE(9)
#else
rsp_status launch_f32_unit7(int S, const LaunchCtx &c, const F32Args &a, cudaStream_t st, const char **err) {
  switch (S) {
    RSP_CASE(10) RSP_CASE(11) RSP_CASE(12) RSP_CASE(13) RSP_CASE(14)
#endif
  default: *err = "transform size must be 2..16384"; return RSP_EUNSUPPORTED;
  }
}
#endif

}  // namespace rsp
