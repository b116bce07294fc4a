// rsp_fast.cu -- specialised kernels for the chain configurations the benchmarks run.
#include "rsp_fast.cuh"

namespace rsp {
bool fast_chain_supported(const Plan &, const FxArgs &) { return false; }
rsp_status launch_fast_chain(const Plan &, const LaunchCtx &, const FxArgs &, cudaStream_t, const char **err) {
  *err = "no fast variant for this configuration";
  return RSP_EUNSUPPORTED;
}
}  // namespace rsp
