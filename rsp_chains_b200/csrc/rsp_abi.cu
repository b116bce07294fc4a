// rsp_abi.cu -- the C ABI of librsp_b200.so (include/rsp_chains.h): handle, register files, kernel dispatch and the
// host-buffer pipeline.  CUDA only: without a device every compute entry point fails with RSP_ECUDA.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <string>
#include <vector>

#include "../../include/rsp_chains.h"
#include "rsp_host.hpp"
#include "rsp_launch.hpp"
#include "rsp_fast.cuh"
#include "rsp_doppler.cuh"

using namespace rsp;

static thread_local std::string g_last_error;
static rsp_status fail(rsp_status st, const std::string &msg) { g_last_error = msg; return st; }
extern "C" void rsp_set_last_error_(const char *msg) { g_last_error = msg ? msg : ""; }   // for the other units of the library (rsp_nccl.cu)
#define CUDA_TRY(expr)                                                                                       \
  do {                                                                                                       \
    cudaError_t e__ = (expr);                                                                                \
    if (e__ != cudaSuccess) return fail(RSP_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));     \
  } while (0)

static int env_int_early(const char *name) { const char *v = std::getenv(name); return (v && *v && *v != '0') ? 1 : 0; }
// Every entry point runs on the handle's device and leaves the caller's current device as it found it (a single-process program
// with several GPUs, e.g. torch, keeps its own notion of the current device).
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
// RSP_DEBUG_ERRORS=1: report (stderr) a CUDA error left pending by an earlier call as soon as the next entry point sees it
static void debug_stale(const char *where) {
  static const int on = env_int_early("RSP_DEBUG_ERRORS");
  if (!on) return;
  const cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) std::fprintf(stderr, "[rsp] pending CUDA error at entry of %s: %s\n", where, cudaGetErrorString(e));
}
#define RSP_ON_DEVICE(c)                                                                                     \
  debug_stale(__func__);                                                                                     \
  DeviceGuard dev_guard__((c)->device);                                                                      \
  if (dev_guard__.err != cudaSuccess) return fail(RSP_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(dev_guard__.err))

static const int N_PIPE = 6;   // most chunks in flight in rsp_run_host (pipe_depth() of them are used)
// tuning knobs of the host pipeline (defaults measured on B200 / PCIe Gen5, scripts/e2e_sweep.py)
static int env_int(const char *name, int dflt, int lo, int hi) {
  const char *v = std::getenv(name);
  if (!v || !*v) return dflt;
  const long x = std::strtol(v, nullptr, 10);
  return x < lo ? lo : (x > hi ? hi : (int)x);
}
static int pipe_depth() { static const int d = env_int("RSP_PIPE_DEPTH", 3, 1, N_PIPE); return d; }
static bool pipe_ramp() { static const int r = env_int("RSP_PIPE_RAMP", 1, 0, 1); return r != 0; }
// chunk size of rsp_run_host: RSP_PIPE_CHUNK_MIB when set, else 1/32 of the batch within 16 .. 64 MiB (measured on B200 / PCIe Gen5:
// 512 MiB batches 16 MiB chunks 42.7 GB/s each way (8 MiB 39.9), 2 GiB batches 16 / 32 / 64 / 128 MiB 43.2 / 45.3 / 45.7 / 45.4,
// profiles/r01_e_e2e_pipeline_sweep.txt, profiles/r02_l_e2e_sweep.txt)
static size_t pipe_chunk_bytes(size_t batch_bytes) {
  static const int m = env_int("RSP_PIPE_CHUNK_MIB", 0, 0, 1024);
  if (m > 0) return (size_t)m << 20;
  return std::min<size_t>((size_t)64 << 20, std::max<size_t>((size_t)16 << 20, batch_bytes / 32));
}

struct rsp_chain {
  rsp_chain_params p;
  Regs r;
  Tables t;
  int device = 0, sm_count = 0;
  uint32_t *d_tw = nullptr;
  int32_t *d_tw2 = nullptr, *d_tww = nullptr;
  uint32_t *d_ovf = nullptr;           // per-stage overflow status of the last call (FFTParams.overflowReg)
  F2 *d_twf = nullptr;
  float *d_twf4 = nullptr;
  int32_t *d_log_tab = nullptr, *d_nco_tab = nullptr, *d_win = nullptr, *d_win4 = nullptr;
  uint32_t *d_nco_full = nullptr;
  float *d_winf = nullptr;
  GenSegment *d_segs = nullptr;
  size_t segs_cap = 0;
  uint16_t *d_gen_phase = nullptr;     // GenPlan::phase_tab on the device (the generator fused into the chain kernel)
  size_t gen_phase_cap = 0;
  uint32_t *d_gen_scratch = nullptr;   // NCO beats between generator and chain when the caller does not want them
  size_t gen_scratch_cap = 0;
  bool gen_dirty = true;               // PLFG registers / RAM changed since the segment list was uploaded
  bool win_dirty = false;              // window RAM of FFTBlockWithWindowing written since the coefficient arrays were uploaded
  GenPlan gen_plan;
  uint64_t launches = 0;
  std::string variant;
  // host pipeline
  cudaStream_t pipe_stream[N_PIPE] = {};
  void *pipe_in[N_PIPE] = {}, *pipe_out[N_PIPE] = {};
  uint32_t *pipe_mask[N_PIPE] = {};
  void *stage_in[N_PIPE] = {}, *stage_out[N_PIPE] = {};   // pinned staging for pageable callers
  size_t pipe_rows_cap = 0, pipe_row_in = 0, pipe_row_out = 0, pipe_row_mask = 0;
  rsp_detection *d_det_list = nullptr;
  uint32_t *d_det_count = nullptr;
  uint32_t det_cap = 0;
};

// ---------------------------------------------------------------------------------------------------------------------
// small kernels that live in this translation unit
namespace rsp {
// LogMagMux alone: one beat per thread, grid-stride (AXI4LogMagMuxBlock)
__global__ void __launch_bounds__(256) mag_fixed_kernel(const FxArgs a, unsigned long long n_beats) {
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_beats; i += stride) {
    const uint32_t wd = ldg(&a.in[i]);
    const int re = wrapw((int)(wd >> 16), a.mag_w), im = wrapw((int)(wd & 0xFFFFu), a.mag_w);
    int ore;
    const int v = mag_one(a, re, im, &ore);
    a.out[i] = pack_iq(ore, v);
  }
}

__global__ void __launch_bounds__(256) mag_f32_kernel(const F32Args a, unsigned long long n_beats) {
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_beats; i += stride) {
    const F2 v = a.in[i];
    a.out[i] = mag_f32(a.sel, v.x, v.y);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
#ifndef RSP_GEN_RUN
#define RSP_GEN_RUN 8
#endif
// PLFG -> NCO: a thread produces runs of RSP_GEN_RUN consecutive samples (closed form for the first, recurrences for the rest, rsp_device_gen.cuh)
// and stores them as 16-byte words when the buffers allow: a warp writes 1 KB contiguous per run (runs of 16 measured slower: 280 vs 328 Gsamples/s)
__global__ void __launch_bounds__(256) generator_kernel(const GenArgs g, unsigned long long n_samples, uint32_t *nco_out, int32_t *plfg_out) {
  constexpr int V = RSP_GEN_RUN;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * V;
  const bool vec = ((reinterpret_cast<uintptr_t>(nco_out) | reinterpret_cast<uintptr_t>(plfg_out)) & 15u) == 0;
  for (unsigned long long k0 = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * V; k0 < n_samples; k0 += stride) {
    uint32_t beat[V];
    int val[V];
    GenState st;
    gen_start(g, k0, st);
    beat[0] = gen_beat(g, st, &val[0]);
#pragma unroll
    for (int j = 1; j < V; ++j) {
      gen_next(g, k0 + j, st);          // past the end of the request: computed, never stored
      beat[j] = gen_beat(g, st, &val[j]);
    }
    if (vec && k0 + V <= n_samples) {
      if (nco_out) {
#pragma unroll
        for (int j = 0; j < V; j += 4) *reinterpret_cast<uint4 *>(nco_out + k0 + j) = make_uint4(beat[j], beat[j + 1], beat[j + 2], beat[j + 3]);
      }
      if (plfg_out) {
#pragma unroll
        for (int j = 0; j < V; j += 4) *reinterpret_cast<int4 *>(plfg_out + k0 + j) = make_int4(val[j], val[j + 1], val[j + 2], val[j + 3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < V; ++j)
        if (k0 + j < n_samples) {
          if (nco_out) nco_out[k0 + j] = beat[j];
          if (plfg_out) plfg_out[k0 + j] = val[j];
        }
    }
  }
}

// The NCO beats alone from the period's phase table (GenPlan::phase_tab, the table the fused chain kernel reads): sample k of the stream has
// the phase (phase[k mod period] + (k div period) * pp) mod 2^phaseWidth and its beat is full[phase] -- two loads and a handful of integer
// instructions per sample instead of the closed form's segment search and recurrences.  Taken when the caller does not ask for the PLFG
// words and the tables exist (phaseWidth <= 12, period up to 2^20 samples).
__global__ void __launch_bounds__(256) generator_table_kernel(const GenArgs g, const uint16_t *__restrict__ phase, uint32_t period, uint32_t pp, uint32_t mask,
                                                              unsigned long long n_samples, uint32_t *__restrict__ nco_out) {
  constexpr int V = RSP_GEN_RUN;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * V;
  const bool vec = (reinterpret_cast<uintptr_t>(nco_out) & 15u) == 0;
  for (unsigned long long k0 = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * V; k0 < n_samples; k0 += stride) {
    unsigned long long per64 = k0 / period;
    uint32_t q = (uint32_t)(k0 - per64 * period), per = (uint32_t)per64;   // (per * pp is taken mod 2^32, which contains 2^phaseWidth)
    uint32_t beat[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const uint32_t ph = ((uint32_t)__ldg(&phase[q]) + per * pp) & mask;
      const uint32_t b = __ldg(&g.nco_full[ph]);
      beat[j] = g.mult_on ? nco_multiply(g, b) : b;
      if (++q == period) { q = 0; ++per; }
    }
    if (vec && k0 + V <= n_samples) {
#pragma unroll
      for (int j = 0; j < V; j += 4) *reinterpret_cast<uint4 *>(nco_out + k0 + j) = make_uint4(beat[j], beat[j + 1], beat[j + 2], beat[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < V; ++j) if (k0 + j < n_samples) nco_out[k0 + j] = beat[j];
    }
  }
}

}  // namespace rsp

// ---------------------------------------------------------------------------------------------------------------------
// launch helpers
static rsp_status launch_fixed(rsp_chain *c, const Plan &pl, const FxArgs &a, cudaStream_t st) {
  if (a.n_rows == 0) return RSP_OK;
  const LaunchCtx ctx = {c->device, c->sm_count};
  const char *err = "";
  const rsp_status s = fast_chain_supported(pl, a) ? launch_fast_chain(pl, ctx, a, st, &err)
                       : fast_fft_supported(pl, a) ? launch_fast_fft(pl, ctx, a, st, &err)
                                                   : launch_fixed_inst(pl.mode, pl.S, ctx, a, st, &err);
  if (s != RSP_OK) return fail(s, err);
  ++c->launches;
  return RSP_OK;
}

static rsp_status launch_f32(rsp_chain *c, const Plan &pl, const F32Args &a, cudaStream_t st) {
  if (a.n_rows == 0) return RSP_OK;
  const LaunchCtx ctx = {c->device, c->sm_count};
  const char *err = "";
  const rsp_status s = fast_f32_chain_supported(pl, a) ? launch_fast_f32_chain(pl, ctx, a, reinterpret_cast<const float4 *>(c->d_twf4), st, &err)
                                                       : launch_f32_inst(pl.S, ctx, a, st, &err);
  if (s != RSP_OK) return fail(s, err);
  ++c->launches;
  return RSP_OK;
}

static unsigned ew_grid(const rsp_chain *c, unsigned long long n) {
  const unsigned long long want = (n + 255) / 256;
  return (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>(want, (unsigned long long)c->sm_count * 16));
}

// ---------------------------------------------------------------------------------------------------------------------
// life cycle
extern "C" void rsp_chain_params_default(rsp_chain_params *p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->arith = RSP_ARITH_FIXED;
  p->hasGenerator = 1;
  // RspChainVanillaApp (src/main/scala/RspChain.scala:83-147)
  p->plfg = {4, 8, 8, 4, 4, 8, 16, 0};
  p->nco.tableSize = 128; p->nco.tableWidth = 16; p->nco.phaseWidth = 9; p->nco.phaseAccEnable = 1;
  rsp_fft_params &f = p->fft;
  f.numPoints = 1024; f.dataWidth = 16; f.binPoint = 0; f.twiddleWidth = 16; f.trimType = RSP_TRIM_ROUND_HALF_UP; f.use4Muls = 0;
  f.useBitReverse = 1; f.fftDir = 1; f.runTime = 1; f.keepMSBorLSB = 0x3FF; f.decimType = 0; f.sdfRadix = 4; f.expandLogic = 0;
  f.keepMSBorLSBReg = 0; f.fftDirReg = 0; f.windowFunc = RSP_WINDOW_NONE; f.windowWidth = 16;
  p->mag = {16, 0, 16, 9, 9};
  rsp_cfar_params &c = p->cfar;
  c.inWidth = 16; c.inBinPoint = 0; c.thresholdWidth = 16; c.thresholdBinPoint = 0; c.scalerWidth = 16; c.scalerBinPoint = 0;
  c.CFARAlgorithm = RSP_CFAR_CA; c.includeCASH = 0; c.minSubWindowSize = 0; c.leadLaggWindowSize = 64; c.sendCut = 0;
  c.guardWindowSize = 4; c.fftSize = 1024;
}

static void free_pipe(rsp_chain *c) {
  for (int i = 0; i < N_PIPE; ++i) {
    if (c->pipe_in[i]) cudaFree(c->pipe_in[i]);
    if (c->pipe_out[i]) cudaFree(c->pipe_out[i]);
    if (c->pipe_mask[i]) cudaFree(c->pipe_mask[i]);
    if (c->stage_in[i]) cudaFreeHost(c->stage_in[i]);
    if (c->stage_out[i]) cudaFreeHost(c->stage_out[i]);
    c->pipe_in[i] = c->pipe_out[i] = nullptr; c->pipe_mask[i] = nullptr; c->stage_in[i] = c->stage_out[i] = nullptr;
  }
  c->pipe_rows_cap = 0;
}

extern "C" void rsp_chain_destroy(rsp_chain *c) {
  if (!c) return;
  debug_stale("rsp_chain_destroy");
  DeviceGuard guard(c->device);
  free_pipe(c);
  for (int i = 0; i < N_PIPE; ++i) if (c->pipe_stream[i]) cudaStreamDestroy(c->pipe_stream[i]);
#define RSP_FREE(ptr) do { const cudaError_t e__ = cudaFree(ptr); if (e__ != cudaSuccess && env_int_early("RSP_DEBUG_ERRORS")) std::fprintf(stderr, "[rsp] cudaFree(" #ptr "): %s\n", cudaGetErrorString(e__)); } while (0)
  RSP_FREE(c->d_tw); RSP_FREE(c->d_tw2); RSP_FREE(c->d_tww); RSP_FREE(c->d_ovf); RSP_FREE(c->d_twf); RSP_FREE(c->d_twf4); RSP_FREE(c->d_win); RSP_FREE(c->d_win4);
  RSP_FREE(c->d_winf); RSP_FREE(c->d_log_tab); RSP_FREE(c->d_nco_tab); RSP_FREE(c->d_nco_full); RSP_FREE(c->d_segs);
  RSP_FREE(c->d_gen_scratch); RSP_FREE(c->d_gen_phase); RSP_FREE(c->d_det_list); RSP_FREE(c->d_det_count);
#undef RSP_FREE
  debug_stale("rsp_chain_destroy (end)");
  delete c;
}

template <typename T>
static rsp_status upload(const std::vector<T> &v, T **d) {
  CUDA_TRY(cudaMalloc((void **)d, std::max<size_t>(1, v.size()) * sizeof(T)));
  if (!v.empty()) CUDA_TRY(cudaMemcpy(*d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return RSP_OK;
}

extern "C" rsp_status rsp_chain_create(const rsp_chain_params *params, int device, rsp_chain **out) {
  if (!params || !out) return fail(RSP_EINVAL, "null argument");
  *out = nullptr;
  std::string err;
  const rsp_status vs = validate_params(*params, err);
  if (vs != RSP_OK) return fail(vs, err);
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1)
    return fail(RSP_ECUDA, "no CUDA device: librsp_b200 has no CPU path");
  if (device < 0 || device >= n_dev) return fail(RSP_EINVAL, "device index out of range");
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess) return fail(RSP_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(guard.err));
  rsp_chain *c = new (std::nothrow) rsp_chain();
  if (!c) return fail(RSP_ENOMEM, "out of host memory");
  c->p = *params;
  c->device = device;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete c; return fail(RSP_ECUDA, "cudaGetDeviceProperties failed"); }
  c->sm_count = prop.multiProcessorCount;
  regs_reset(c->p, c->r);
  build_tables(c->p, c->t);
  build_window(c->p, c->t);
  rsp_status s = upload(c->t.tw, &c->d_tw);
  if (s == RSP_OK) s = upload(c->t.tw2, &c->d_tw2);
  if (s == RSP_OK && (fft_is_wide(c->p.fft) || c->p.fft.binPoint >= 2)) s = upload(c->t.tww, &c->d_tww);   // raw twiddles: the wide block and the binPoint >= 2 fast kernels
  if (s == RSP_OK && c->p.fft.overflowReg) {
    if (cudaMalloc((void **)&c->d_ovf, sizeof(uint32_t)) != cudaSuccess || cudaMemset(c->d_ovf, 0, sizeof(uint32_t)) != cudaSuccess) s = fail(RSP_ECUDA, "allocating the overflow status word failed");
  }
  if (s == RSP_OK) s = upload(c->t.twf, &c->d_twf);
  if (s == RSP_OK) s = upload(c->t.twf4, &c->d_twf4);
  if (s == RSP_OK) s = upload(c->t.win, &c->d_win);
  if (s == RSP_OK) s = upload(c->t.win4, &c->d_win4);
  if (s == RSP_OK) s = upload(c->t.winf, &c->d_winf);
  if (s == RSP_OK) s = upload(c->t.log_tab, &c->d_log_tab);
  if (s == RSP_OK) s = upload(c->t.nco_tab, &c->d_nco_tab);
  if (s == RSP_OK && !c->t.nco_full.empty()) s = upload(c->t.nco_full, &c->d_nco_full);
  if (s != RSP_OK) { rsp_chain_destroy(c); return s; }
  *out = c;
  return RSP_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// registers
// FFTBlockWithWindowing has doWindowing at word 2 and the rest of AXI4FFTBlock's file one word further (FFTBlockWithWindowing.scala:134-152)
static bool fftw_translate(const rsp_chain *c, rsp_block blk, uint32_t *idx, bool *is_dowin) {
  *is_dowin = false;
  if (blk != RSP_BLK_FFT || !c->p.fft.windowRam) return true;
  if (*idx == RSP_FFTW_REG_DOWINDOWING) { *is_dowin = true; return true; }
  if (*idx > RSP_FFTW_REG_DOWINDOWING) *idx -= 1;
  return true;
}

extern "C" rsp_status rsp_reg_write(rsp_chain *c, rsp_block blk, uint32_t idx, uint32_t value) {
  if (!c) return fail(RSP_EINVAL, "null handle");
  bool dowin = false;
  fftw_translate(c, blk, &idx, &dowin);
  if (dowin) { c->r.fft_dowin = value & 1u; return RSP_OK; }
  bool ro = false;
  const int w = reg_width(c->p, blk, idx, &ro);
  if (blk == RSP_BLK_CFAR && idx == RSP_CFAR_REG_THRESHOLDSCALER_F32) {
    std::memcpy(&c->r.scaler_f32, &value, 4);
    return RSP_OK;
  }
  if (w < 0) return fail(RSP_EINVAL, "no such register");
  if (ro) return RSP_OK;   // writes to read-only fields are ignored by the regmap
  uint32_t *slot = reg_slot(c->r, blk, idx);
  if (!slot) return fail(RSP_EINVAL, "no such register");
  *slot = w >= 32 ? value : (value & ((1u << w) - 1u));   // RegField(width): upper bits are dropped
  if (blk == RSP_BLK_PLFG) c->gen_dirty = true;
  return RSP_OK;
}

extern "C" rsp_status rsp_reg_read(rsp_chain *c, rsp_block blk, uint32_t idx, uint32_t *value) {
  if (!c || !value) return fail(RSP_EINVAL, "null argument");
  if (blk == RSP_BLK_CFAR && idx == RSP_CFAR_REG_THRESHOLDSCALER_F32) { std::memcpy(value, &c->r.scaler_f32, 4); return RSP_OK; }
  bool dowin = false;
  fftw_translate(c, blk, &idx, &dowin);
  if (dowin) { *value = c->r.fft_dowin; return RSP_OK; }
  bool ro = false;
  if (reg_width(c->p, blk, idx, &ro) < 0) return fail(RSP_EINVAL, "no such register");
  uint32_t *slot = reg_slot(c->r, blk, idx);
  if (!slot) return fail(RSP_EINVAL, "no such register");
  *value = *slot;   // busy reads 0: calls are synchronous with respect to the registers
  if (blk == RSP_BLK_FFT && idx == RSP_FFT_REG_OVERFLOW && c->d_ovf != nullptr) {
    // the status of the most recent call, through the reference's log2Ceil(numStages)-bit field (FFTBlock.scala:67,82,97)
    RSP_ON_DEVICE(c);
    uint32_t m = 0;
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpy(&m, c->d_ovf, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    const int w = reg_width(c->p, blk, idx, &ro);
    *value = m & ((1u << w) - 1u);
  }
  return RSP_OK;
}

extern "C" rsp_status rsp_plfg_ram_write(rsp_chain *c, uint32_t idx, uint32_t value) {
  if (!c) return fail(RSP_EINVAL, "null handle");
  if (!c->p.hasGenerator) return fail(RSP_EINVAL, "this chain has no PLFG");
  if (idx >= c->r.plfg_ram.size()) return fail(RSP_EINVAL, "PLFG RAM index out of range");
  const int bits = c->p.plfg.outputWidthInt + c->p.plfg.maxNumOfSamplesWidth + 2;   // protoRAM width (PLFGDspBlockMem.scala:108-117)
  c->r.plfg_ram[idx] = bits >= 32 ? value : (value & ((1u << bits) - 1u));
  c->gen_dirty = true;
  return RSP_OK;
}

extern "C" rsp_status rsp_fft_window_user(rsp_chain *c, const double *coeffs, uint32_t n) {
  if (!c || !coeffs) return fail(RSP_EINVAL, "null argument");
  if (c->p.fft.windowFunc != RSP_WINDOW_USER) return fail(RSP_EINVAL, "the chain was not created with windowFunc = RSP_WINDOW_USER");
  if (n != (uint32_t)c->p.fft.numPoints) return fail(RSP_EINVAL, "Length of specified window function is not the same as fft size");
  RSP_ON_DEVICE(c);
  CUDA_TRY(cudaDeviceSynchronize());   // a kernel in flight may still read the old ROM
  c->t.win_user.assign(coeffs, coeffs + n);
  build_window(c->p, c->t);
  CUDA_TRY(cudaMemcpy(c->d_win, c->t.win.data(), c->t.win.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->d_win4, c->t.win4.data(), c->t.win4.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->d_winf, c->t.winf.data(), c->t.winf.size() * sizeof(float), cudaMemcpyHostToDevice));
  return RSP_OK;
}

extern "C" rsp_status rsp_fft_window_ram_write(rsp_chain *c, uint32_t idx, uint32_t value) {
  if (!c) return fail(RSP_EINVAL, "null handle");
  if (!c->p.fft.windowRam) return fail(RSP_EINVAL, "this chain's FFT block has no window RAM (windowRam = 0)");
  // w_addr keeps log2Ceil(numPoints) address bits (FFTBlockWithWindowing.scala:54): higher word indices alias
  idx &= (uint32_t)c->p.fft.numPoints - 1u;
  const int w = c->p.fft.windowWidth;
  if (c->t.win_ram.size() != (size_t)c->p.fft.numPoints) c->t.win_ram.assign((size_t)c->p.fft.numPoints, 0);
  c->t.win_ram[idx] = (int32_t)wrap_host((long long)value, w);   // wdata.asTypeOf(protoWin) (:51)
  c->win_dirty = true;
  return RSP_OK;
}

// coefficient arrays on the device follow the window RAM lazily: uploaded by the first run call after a write
static rsp_status sync_window(rsp_chain *c) {
  if (!c->win_dirty) return RSP_OK;
  CUDA_TRY(cudaDeviceSynchronize());   // a kernel in flight may still read the old coefficients
  build_window(c->p, c->t);
  CUDA_TRY(cudaMemcpy(c->d_win, c->t.win.data(), c->t.win.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->d_win4, c->t.win4.data(), c->t.win4.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->d_winf, c->t.winf.data(), c->t.winf.size() * sizeof(float), cudaMemcpyHostToDevice));
  c->win_dirty = false;
  return RSP_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// planning glue
static void bind_fixed(rsp_chain *c, FxArgs &a) { a.tw = c->d_tw; a.tw2 = c->d_tw2; a.tww = c->d_tww; a.log_tab = c->d_log_tab; a.win = c->d_win; a.win4 = c->d_win4; }
static void bind_f32(rsp_chain *c, F32Args &a) { a.tw = c->d_twf; a.win = c->d_winf; }
static void bind_dets(const rsp_dets *d, Detection **list, uint32_t **count, uint32_t *cap) {
  if (d && d->list && d->count) { *list = reinterpret_cast<Detection *>(d->list); *count = d->count; *cap = d->capacity; }
  else { *list = nullptr; *count = nullptr; *cap = 0; }
}

struct FusedGen { const uint16_t *phase; const uint32_t *full; uint32_t period, pp, mask; };
static rsp_status run_stages(rsp_chain *c, int first, int last, const void *d_in, uint32_t n_rows, void *d_out, uint32_t *d_peak_mask,
                             const rsp_dets *dets, cudaStream_t st, uint32_t row_base = 0, const FusedGen *fg = nullptr) {
  std::string err;
  Plan pl;
  { const rsp_status ws = sync_window(c); if (ws != RSP_OK) return ws; }
  if (c->p.arith == RSP_ARITH_FIXED) {
    FxArgs a;
    const rsp_status s = plan_fixed(c->p, c->r, c->t, first, last, a, pl, err);
    if (s != RSP_OK) return fail(s, err);
    bind_fixed(c, a);
    if (first == STAGE_FFT && c->d_ovf != nullptr) {
      a.ovf_flags = c->d_ovf;
      if (row_base == 0) CUDA_TRY(cudaMemsetAsync(c->d_ovf, 0, sizeof(uint32_t), st));   // the status is that of this call (all its chunks)
    }
    a.in = static_cast<const uint32_t *>(d_in); a.out = static_cast<uint32_t *>(d_out); a.n_rows = n_rows; a.row_base = row_base;
    bind_dets(dets, &a.det_list, &a.det_count, &a.det_cap);
    if (!a.out && !(last == STAGE_CFAR && a.det_list)) return fail(RSP_EINVAL, "no output requested");
    if (fg != nullptr) {   // the generator fused into the first pass: only the fast CA kernel reads these
      if (!(fast_chain_supported(pl, a) && a.core == CORE_MEM)) return fail(RSP_EUNSUPPORTED, "no fused-generator kernel for this configuration");
      a.gen_phase = fg->phase; a.gen_full = fg->full; a.gen_period = fg->period; a.gen_pp = fg->pp; a.gen_mask = fg->mask; a.gen_k0 = 0;
    }
    return launch_fixed(c, pl, a, st);
  }
  F32Args a;
  const rsp_status s = plan_f32(c->p, c->r, c->t, first, last, a, pl, err);
  if (s != RSP_OK) return fail(s, err);
  bind_f32(c, a);
  if (first == STAGE_CFAR) a.in_mag = static_cast<const float *>(d_in); else a.in = static_cast<const F2 *>(d_in);
  if (last == STAGE_FFT) a.out_c = static_cast<F2 *>(d_out); else a.out = static_cast<float *>(d_out);
  a.peak_mask = d_peak_mask; a.n_rows = n_rows; a.row_base = row_base;
  bind_dets(dets, &a.det_list, &a.det_count, &a.det_cap);
  if (!d_out && !(last == STAGE_CFAR && (a.det_list || a.peak_mask))) return fail(RSP_EINVAL, "no output requested");
  return launch_f32(c, pl, a, st);
}

static rsp_status rows_of(uint32_t n_chirps, uint32_t n_rx, uint32_t *rows) {
  const unsigned long long r = (unsigned long long)n_chirps * (unsigned long long)n_rx;
  if (r > 0x7FFFFFFFull) return fail(RSP_EINVAL, "n_chirps * n_rx too large");
  *rows = (uint32_t)r;
  return RSP_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// device entry points
extern "C" rsp_status rsp_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                              const rsp_dets *dets, void *stream) {
  if (!c || !d_in) return fail(RSP_EINVAL, "null argument");
  uint32_t rows;
  rsp_status s = rows_of(n_chirps, n_rx, &rows);
  if (s != RSP_OK) return s;
  RSP_ON_DEVICE(c);
  return run_stages(c, STAGE_FFT, STAGE_CFAR, d_in, rows, d_out, d_peak_mask, dets, static_cast<cudaStream_t>(stream));
}

extern "C" rsp_status rsp_fft_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, void *d_out, void *stream) {
  if (!c || !d_in || !d_out) return fail(RSP_EINVAL, "null argument");
  RSP_ON_DEVICE(c);
  if (c->p.arith == RSP_ARITH_FIXED && fft_is_wide(c->p.fft)) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    WideArgs a;
    std::string err;
    const rsp_status s = plan_wide(c->p, c->r, c->t, a, err);
    if (s != RSP_OK) return fail(s, err);
    if (n_chirps == 0) return RSP_OK;
    { const rsp_status ws = sync_window(c); if (ws != RSP_OK) return ws; }
    a.in = static_cast<const unsigned long long *>(d_in); a.out = static_cast<unsigned long long *>(d_out); a.n_rows = n_chirps;
    a.tw = reinterpret_cast<const int2 *>(c->d_tww); a.win = c->d_win;
    a.ovf_flags = c->d_ovf;
    if (c->d_ovf != nullptr) CUDA_TRY(cudaMemsetAsync(c->d_ovf, 0, sizeof(uint32_t), st));
    const LaunchCtx ctx = {c->device, c->sm_count};
    const char *e = "";
    const rsp_status ls = launch_fft_wide(ctx, a, st, &e);
    if (ls != RSP_OK) return fail(ls, e);
    ++c->launches;
    return RSP_OK;
  }
  return run_stages(c, STAGE_FFT, STAGE_FFT, d_in, n_chirps, d_out, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" rsp_status rsp_cfar_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, void *d_out, uint32_t *d_peak_mask,
                                   const rsp_dets *dets, void *stream) {
  if (!c || !d_in) return fail(RSP_EINVAL, "null argument");
  RSP_ON_DEVICE(c);
  return run_stages(c, STAGE_CFAR, STAGE_CFAR, d_in, n_chirps, d_out, d_peak_mask, dets, static_cast<cudaStream_t>(stream));
}

extern "C" rsp_status rsp_mag_run(rsp_chain *c, const void *d_in, uint64_t n_beats, void *d_out, void *stream) {
  if (!c || !d_in || !d_out) return fail(RSP_EINVAL, "null argument");
  RSP_ON_DEVICE(c);
  if (n_beats == 0) return RSP_OK;
  std::string err;
  Plan pl;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (c->p.arith == RSP_ARITH_FIXED) {
    FxArgs a;
    const rsp_status s = plan_fixed(c->p, c->r, c->t, STAGE_MAG, STAGE_MAG, a, pl, err);
    if (s != RSP_OK) return fail(s, err);
    bind_fixed(c, a);
    a.in = static_cast<const uint32_t *>(d_in); a.out = static_cast<uint32_t *>(d_out);
    mag_fixed_kernel<<<ew_grid(c, n_beats), 256, 0, st>>>(a, n_beats);
  } else {
    F32Args a;
    const rsp_status s = plan_f32(c->p, c->r, c->t, STAGE_MAG, STAGE_MAG, a, pl, err);
    if (s != RSP_OK) return fail(s, err);
    a.in = static_cast<const F2 *>(d_in); a.out = static_cast<float *>(d_out);
    mag_f32_kernel<<<ew_grid(c, n_beats), 256, 0, st>>>(a, n_beats);
  }
  CUDA_TRY(cudaGetLastError());
  ++c->launches;
  return RSP_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// slow-time dimension (kernels in rsp_doppler.cuh)
template <typename T, int TILE>
static void launch_corner_turn(const rsp_chain *c, const void *d_in, void *d_out, uint32_t n_rows, uint32_t n_cols, uint32_t batch, cudaStream_t st) {
  const uint32_t tr = (n_rows + TILE - 1) / TILE, tc = (n_cols + TILE - 1) / TILE;
  const unsigned long long n_tiles = (unsigned long long)tr * tc * batch;
  const unsigned grid = (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>(n_tiles, (unsigned long long)c->sm_count * 8));
  corner_turn_kernel<T, TILE><<<grid, 256, 0, st>>>(static_cast<const T *>(d_in), static_cast<T *>(d_out), n_rows, n_cols, tr, tc, n_tiles);
}

extern "C" rsp_status rsp_corner_turn(rsp_chain *c, const void *d_in, void *d_out, uint32_t n_rows, uint32_t n_cols, uint32_t batch,
                                      uint32_t elem_bytes, void *stream) {
  if (!c || !d_in || !d_out) return fail(RSP_EINVAL, "null argument");
  if (d_in == d_out) return fail(RSP_EINVAL, "rsp_corner_turn is out of place");
  if (elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return fail(RSP_EINVAL, "rsp_corner_turn: elem_bytes must be 2, 4 or 8");
  RSP_ON_DEVICE(c);
  if (n_rows == 0 || n_cols == 0 || batch == 0) return RSP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (elem_bytes == 2) launch_corner_turn<uint16_t, 64>(c, d_in, d_out, n_rows, n_cols, batch, st);
  else if (elem_bytes == 4) launch_corner_turn<uint32_t, 32>(c, d_in, d_out, n_rows, n_cols, batch, st);
  else launch_corner_turn<unsigned long long, 32>(c, d_in, d_out, n_rows, n_cols, batch, st);
  CUDA_TRY(cudaGetLastError());
  ++c->launches;
  return RSP_OK;
}

extern "C" rsp_status rsp_mag_combine_run(rsp_chain *c, const void *d_in, uint64_t n_beats, uint32_t n_rx, void *d_out, void *stream) {
  if (!c || !d_in || !d_out) return fail(RSP_EINVAL, "null argument");
  if (n_rx == 0 || (n_rx & (n_rx - 1)) || n_rx > 64) return fail(RSP_EINVAL, "rsp_mag_combine_run: n_rx must be a power of two <= 64 (the mean is a shift)");
  RSP_ON_DEVICE(c);
  if (n_beats == 0) return RSP_OK;
  std::string err;
  Plan pl;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int shift = 0;
  while ((1u << shift) < n_rx) ++shift;
  if (c->p.arith == RSP_ARITH_FIXED) {
    FxArgs a;
    const rsp_status s = plan_fixed(c->p, c->r, c->t, STAGE_MAG, STAGE_MAG, a, pl, err);
    if (s != RSP_OK) return fail(s, err);
    if (a.sel == 0) return fail(RSP_EINVAL, "rsp_mag_combine_run: the magnitude selector is 0 (bypass): there is no magnitude to combine");
    bind_fixed(c, a);
    a.in = static_cast<const uint32_t *>(d_in); a.out = static_cast<uint32_t *>(d_out);
    mag_combine_fixed_kernel<<<ew_grid(c, n_beats), 256, 0, st>>>(a, n_beats, n_rx, shift);
  } else {
    F32Args a;
    const rsp_status s = plan_f32(c->p, c->r, c->t, STAGE_MAG, STAGE_MAG, a, pl, err);
    if (s != RSP_OK) return fail(s, err);
    if (a.sel == 0) return fail(RSP_EINVAL, "rsp_mag_combine_run: the magnitude selector is 0 (bypass): there is no magnitude to combine");
    a.in = static_cast<const F2 *>(d_in); a.out = static_cast<float *>(d_out);
    mag_combine_f32_kernel<<<ew_grid(c, n_beats), 256, 0, st>>>(a, n_beats, n_rx, 1.0f / (float)n_rx);
  }
  CUDA_TRY(cudaGetLastError());
  ++c->launches;
  return RSP_OK;
}

static rsp_status prepare_generator(rsp_chain *c, uint64_t n_samples, GenArgs &g, cudaStream_t st) {
  if (!c->p.hasGenerator) return fail(RSP_EINVAL, "this chain has no PLFG/NCO (hasGenerator = 0)");
  if (!(c->r.plfg[RSP_PLFG_REG_ENABLE] & 1u)) return fail(RSP_EINVAL, "the PLFG enable register is 0");
  if (c->gen_dirty) {
    std::string err;
    const rsp_status s = plan_generator(c->p, c->r, c->gen_plan, err);
    if (s != RSP_OK) return fail(s, err);
    const size_t n = c->gen_plan.segs.size();
    if (n > c->segs_cap) {
      // the previous list may still be read by a kernel in flight on another stream: drain before freeing
      CUDA_TRY(cudaDeviceSynchronize());
      cudaFree(c->d_segs); c->d_segs = nullptr; c->segs_cap = 0;
      CUDA_TRY(cudaMalloc((void **)&c->d_segs, n * sizeof(GenSegment)));
      c->segs_cap = n;
    }
    const size_t np = c->gen_plan.phase_tab.size();
    if (np > c->gen_phase_cap) {
      CUDA_TRY(cudaDeviceSynchronize());
      cudaFree(c->d_gen_phase); c->d_gen_phase = nullptr; c->gen_phase_cap = 0;
      CUDA_TRY(cudaMalloc((void **)&c->d_gen_phase, np * sizeof(uint16_t)));
      c->gen_phase_cap = np;
    }
    if (np) CUDA_TRY(cudaMemcpyAsync(c->d_gen_phase, c->gen_plan.phase_tab.data(), np * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(c->d_segs, c->gen_plan.segs.data(), n * sizeof(GenSegment), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));   // the host vector may be rebuilt by the next register write
    c->gen_dirty = false;
  }
  if (n_samples > c->gen_plan.total)
    return fail(RSP_EINVAL, "the PLFG schedule (frameNum frames) ends before the requested samples; frameNum = 0 runs forever");
  g.segs = c->d_segs; g.n_segs = (uint32_t)c->gen_plan.segs.size();
  g.period_len = c->gen_plan.period_len; g.period_phase = c->gen_plan.period_phase;
  g.nco_tab = c->d_nco_tab;
  g.nco_full = c->d_nco_full;   // null when phaseWidth > 12
  g.table_size = c->p.nco.tableSize; g.addr_bits = log2_ceil((uint32_t)c->p.nco.tableSize);
  g.table_width = c->p.nco.tableWidth; g.phase_width = c->p.nco.phaseWidth;
  g.out_width = c->p.plfg.outputWidthInt + c->p.plfg.outputWidthFrac;
  g.phase_acc_enable = c->p.nco.phaseAccEnable ? 1 : 0;
  g.rasterized = c->p.nco.rasterizedMode ? 1 : 0;
  g.poff = 0;   // poffType Fixed: no phase-offset port (NCOLazyModuleBlock.scala:246-271)
  g.mult_on = 0; g.mult_f = 0; g.mult_trim = c->p.nco.roundingMode;
  if (c->p.nco.useMultiplier) {
    const int w = c->p.nco.tableWidth;
    // factor = Mux(enableMultiplying, multiplyingFactor, 1.U << (beatBytes * 4 - 2)) read as FixedPoint(w, w - 2) (NCOLazyModuleBlock.scala:354)
    g.mult_f = (int)wrap_host((c->r.nco[0] & 1u) ? (long long)c->r.nco[1] : (1LL << 14), w);
    g.mult_on = (g.mult_f != (1 << (w - 2))) ? 1 : 0;   // times exactly 1.0 is the identity under every trim type
  }
  return RSP_OK;
}

extern "C" rsp_status rsp_generate(rsp_chain *c, uint64_t n_samples, uint32_t *d_nco_out, int32_t *d_plfg_out, void *stream) {
  if (!c || (!d_nco_out && !d_plfg_out)) return fail(RSP_EINVAL, "null argument");
  RSP_ON_DEVICE(c);
  if (n_samples == 0) return RSP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GenArgs g;
  const rsp_status s = prepare_generator(c, n_samples, g, st);
  if (s != RSP_OK) return s;
  static const int table_on = env_int("RSP_GEN_TABLE", 1, 0, 1);   // A/B knob
  if (table_on && d_plfg_out == nullptr && c->d_nco_full != nullptr && !c->gen_plan.phase_tab.empty() && c->gen_plan.period_len <= 0xFFFFFFFFull) {
    const uint32_t pw = (uint32_t)c->p.nco.phaseWidth;
    generator_table_kernel<<<ew_grid(c, (n_samples + RSP_GEN_RUN - 1) / RSP_GEN_RUN), 256, 0, st>>>(g, c->d_gen_phase, (uint32_t)c->gen_plan.period_len,
                                                                                                   c->gen_plan.period_phase, (1u << pw) - 1u, n_samples, d_nco_out);
  } else {
    generator_kernel<<<ew_grid(c, (n_samples + RSP_GEN_RUN - 1) / RSP_GEN_RUN), 256, 0, st>>>(g, n_samples, d_nco_out, d_plfg_out);
  }
  CUDA_TRY(cudaGetLastError());
  ++c->launches;
  return RSP_OK;
}

extern "C" rsp_status rsp_run_from_generator(rsp_chain *c, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                                             const rsp_dets *dets, uint32_t *d_nco_out, void *stream) {
  if (!c) return fail(RSP_EINVAL, "null handle");
  if (c->p.arith != RSP_ARITH_FIXED) return fail(RSP_EUNSUPPORTED, "the PLFG/NCO generator feeds the fixed-point chain only");
  uint32_t rows;
  rsp_status s = rows_of(n_chirps, n_rx, &rows);
  if (s != RSP_OK) return s;
  RSP_ON_DEVICE(c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int S = c->p.fft.runTime ? (int)c->r.fft[RSP_FFT_REG_FFTSIZE] : log2_ceil((uint32_t)c->p.fft.numPoints);
  const uint64_t n_samples = (uint64_t)rows << S;
  // One launch when the caller does not ask for the NCO beats and the configuration runs on the fast CA kernel: the first pass of the
  // chain kernel generates its own beats from the period's phase table (no beats in HBM: 0 B in + 4 B out per sample).
  // RSP_GEN_FUSED=0 keeps the two-launch path (generator -> scratch beats -> chain), which is also what every other case takes.
  static const int fused_on = env_int("RSP_GEN_FUSED", 1, 0, 1);
  if (!d_nco_out && fused_on && c->d_nco_full != nullptr && c->p.hasGenerator && (c->r.plfg[RSP_PLFG_REG_ENABLE] & 1u)) {
    GenArgs g;
    s = prepare_generator(c, n_samples, g, st);
    if (s != RSP_OK) return s;
    if (!c->gen_plan.phase_tab.empty() && !g.mult_on) {   // (the beat table holds unmultiplied beats)
      std::string err; Plan pl; FxArgs a;
      if (plan_fixed(c->p, c->r, c->t, STAGE_FFT, STAGE_CFAR, a, pl, err) == RSP_OK) {
        a.out = static_cast<uint32_t *>(d_out);
        Detection *dl; uint32_t *dc; uint32_t cap;
        bind_dets(dets, &dl, &dc, &cap);
        a.det_list = dl;
        if (fast_chain_supported(pl, a) && a.core == CORE_MEM) {
          const uint32_t pw = (uint32_t)c->p.nco.phaseWidth;
          const FusedGen fg = {c->d_gen_phase, c->d_nco_full, (uint32_t)c->gen_plan.period_len, c->gen_plan.period_phase, (1u << pw) - 1u};
          return run_stages(c, STAGE_FFT, STAGE_CFAR, c->d_nco_full /* not read */, rows, d_out, d_peak_mask, dets, st, 0, &fg);
        }
      }
    }
  }
  uint32_t *beats = d_nco_out;
  if (!beats) {
    if (n_samples > c->gen_scratch_cap) {
      CUDA_TRY(cudaDeviceSynchronize());
      cudaFree(c->d_gen_scratch); c->d_gen_scratch = nullptr; c->gen_scratch_cap = 0;
      CUDA_TRY(cudaMalloc((void **)&c->d_gen_scratch, n_samples * sizeof(uint32_t)));
      c->gen_scratch_cap = n_samples;
    }
    beats = c->d_gen_scratch;
  }
  s = rsp_generate(c, n_samples, beats, nullptr, stream);
  if (s != RSP_OK) return s;
  return run_stages(c, STAGE_FFT, STAGE_CFAR, beats, rows, d_out, d_peak_mask, dets, st);
}

// ---------------------------------------------------------------------------------------------------------------------
// rsp_run_from_generator without any allocation (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).  The two-launch form of the path --
// generator_kernel -> NCO beats in scratch -> chain kernel -- with the scratch beats and the PLFG segment list in file-scope device arrays
// of a fixed size instead of the handle's cudaMalloc'ed buffers.  Every GPU that loads the library holds both arrays (4 MiB + 16 KiB) until
// exit.  All calls on a GPU share them and they keep the last call's contents: nothing has to start at zero (a call overwrites every beat
// and every segment it reads), but two calls that overlap on the GPU (different streams, different handles) race on them.
static __device__ __align__(256) uint32_t g_static_beats[RSP_STATIC_SCRATCH_BEATS];
static __device__ __align__(256) GenSegment g_static_segs[RSP_STATIC_MAX_SEGMENTS];

// In host code the arrays' names are host-side stand-ins, not device addresses: the addresses come from cudaGetSymbolAddress, once per
// device, with that device current (the caller's RSP_ON_DEVICE guard).
static rsp_status static_scratch(int device, uint32_t **beats, GenSegment **segs) {
  static void *beats_of[64] = {}, *segs_of[64] = {};
  if (device < 0 || device >= 64) return fail(RSP_EUNSUPPORTED, "rsp_run_from_generator_static: device ordinal above 63");
  if (beats_of[device] == nullptr) {
    void *pb = nullptr, *ps = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&pb, g_static_beats));
    CUDA_TRY(cudaGetSymbolAddress(&ps, g_static_segs));
    segs_of[device] = ps; beats_of[device] = pb;
  }
  *beats = static_cast<uint32_t *>(beats_of[device]); *segs = static_cast<GenSegment *>(segs_of[device]);
  return RSP_OK;
}

extern "C" rsp_status rsp_run_from_generator_static(rsp_chain *c, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                                                    const rsp_dets *dets, void *stream) {
  if (!c) return fail(RSP_EINVAL, "null handle");
  if (c->p.arith != RSP_ARITH_FIXED) return fail(RSP_EUNSUPPORTED, "the PLFG/NCO generator feeds the fixed-point chain only");
  if (!c->p.hasGenerator) return fail(RSP_EINVAL, "this chain has no PLFG/NCO (hasGenerator = 0)");
  if (!(c->r.plfg[RSP_PLFG_REG_ENABLE] & 1u)) return fail(RSP_EINVAL, "the PLFG enable register is 0");
  uint32_t rows;
  rsp_status s = rows_of(n_chirps, n_rx, &rows);
  if (s != RSP_OK) return s;
  const int S = c->p.fft.runTime ? (int)c->r.fft[RSP_FFT_REG_FFTSIZE] : log2_ceil((uint32_t)c->p.fft.numPoints);
  const uint64_t n_samples = (uint64_t)rows << S;
  if (n_samples > (uint64_t)RSP_STATIC_SCRATCH_BEATS)
    return fail(RSP_EINVAL, "rsp_run_from_generator_static: " + std::to_string(n_samples) + " samples exceed the static scratch of " +
                                std::to_string((unsigned long long)RSP_STATIC_SCRATCH_BEATS) + " beats (RSP_STATIC_SCRATCH_BEATS); use rsp_run_from_generator");
  RSP_ON_DEVICE(c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n_samples == 0) return RSP_OK;
  uint32_t *beats; GenSegment *segs;
  s = static_scratch(c->device, &beats, &segs);
  if (s != RSP_OK) return s;
  // the schedule is planned into a local (host) plan and uploaded on every call: the handle's own device copy belongs to the allocating path
  GenPlan gp;
  std::string err;
  s = plan_generator(c->p, c->r, gp, err);
  if (s != RSP_OK) return fail(s, err);
  if (gp.segs.size() > (size_t)RSP_STATIC_MAX_SEGMENTS)
    return fail(RSP_EINVAL, "rsp_run_from_generator_static: the PLFG schedule has " + std::to_string(gp.segs.size()) + " segments per period, the static list holds " +
                                std::to_string((int)RSP_STATIC_MAX_SEGMENTS) + " (RSP_STATIC_MAX_SEGMENTS); use rsp_run_from_generator");
  if (n_samples > gp.total)
    return fail(RSP_EINVAL, "the PLFG schedule (frameNum frames) ends before the requested samples; frameNum = 0 runs forever");
  CUDA_TRY(cudaMemcpyAsync(segs, gp.segs.data(), gp.segs.size() * sizeof(GenSegment), cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaStreamSynchronize(st));   // gp is a local: the copy has left it before it goes away
  GenArgs g;
  g.segs = segs; g.n_segs = (uint32_t)gp.segs.size();
  g.period_len = gp.period_len; g.period_phase = gp.period_phase;
  g.nco_tab = c->d_nco_tab; g.nco_full = c->d_nco_full;
  g.table_size = c->p.nco.tableSize; g.addr_bits = log2_ceil((uint32_t)c->p.nco.tableSize);
  g.table_width = c->p.nco.tableWidth; g.phase_width = c->p.nco.phaseWidth;
  g.out_width = c->p.plfg.outputWidthInt + c->p.plfg.outputWidthFrac;
  g.phase_acc_enable = c->p.nco.phaseAccEnable ? 1 : 0;
  g.rasterized = c->p.nco.rasterizedMode ? 1 : 0;
  g.poff = 0;
  g.mult_on = 0; g.mult_f = 0; g.mult_trim = c->p.nco.roundingMode;
  if (c->p.nco.useMultiplier) {   // as prepare_generator (NCOLazyModuleBlock.scala:354)
    const int w = c->p.nco.tableWidth;
    g.mult_f = (int)wrap_host((c->r.nco[0] & 1u) ? (long long)c->r.nco[1] : (1LL << 14), w);
    g.mult_on = (g.mult_f != (1 << (w - 2))) ? 1 : 0;
  }
  generator_kernel<<<ew_grid(c, (n_samples + RSP_GEN_RUN - 1) / RSP_GEN_RUN), 256, 0, st>>>(g, n_samples, beats, nullptr);
  CUDA_TRY(cudaGetLastError());
  ++c->launches;
  return run_stages(c, STAGE_FFT, STAGE_CFAR, beats, rows, d_out, d_peak_mask, dets, st);
}

// ---------------------------------------------------------------------------------------------------------------------
// host-buffer pipeline
extern "C" rsp_status rsp_host_alloc(void **p, size_t bytes) {
  if (!p) return fail(RSP_EINVAL, "null argument");
  CUDA_TRY(cudaHostAlloc(p, std::max<size_t>(bytes, 1), cudaHostAllocDefault));
  return RSP_OK;
}
extern "C" void rsp_host_free(void *p) { if (p) cudaFreeHost(p); }

static bool is_pinned(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

static rsp_status ensure_pipe(rsp_chain *c, size_t rows_per_chunk, size_t row_in, size_t row_out, size_t row_mask, bool need_stage_in, bool need_stage_out) {
  for (int i = 0; i < pipe_depth(); ++i)
    if (!c->pipe_stream[i]) CUDA_TRY(cudaStreamCreateWithFlags(&c->pipe_stream[i], cudaStreamNonBlocking));
  const bool fits = c->pipe_rows_cap >= rows_per_chunk && c->pipe_row_in >= row_in && c->pipe_row_out >= row_out && c->pipe_row_mask >= row_mask;
  if (!fits) {
    CUDA_TRY(cudaDeviceSynchronize());
    free_pipe(c);
    for (int i = 0; i < pipe_depth(); ++i) {
      CUDA_TRY(cudaMalloc(&c->pipe_in[i], rows_per_chunk * row_in));
      CUDA_TRY(cudaMalloc(&c->pipe_out[i], std::max<size_t>(1, rows_per_chunk * row_out)));
      CUDA_TRY(cudaMalloc((void **)&c->pipe_mask[i], std::max<size_t>(1, rows_per_chunk * row_mask)));
    }
    c->pipe_rows_cap = rows_per_chunk; c->pipe_row_in = row_in; c->pipe_row_out = row_out; c->pipe_row_mask = row_mask;
  }
  for (int i = 0; i < pipe_depth(); ++i) {
    if (need_stage_in && !c->stage_in[i]) CUDA_TRY(cudaHostAlloc(&c->stage_in[i], c->pipe_rows_cap * c->pipe_row_in, cudaHostAllocDefault));
    if (need_stage_out && !c->stage_out[i]) CUDA_TRY(cudaHostAlloc(&c->stage_out[i], std::max<size_t>(1, c->pipe_rows_cap * std::max(c->pipe_row_out, c->pipe_row_mask)), cudaHostAllocDefault));
  }
  return RSP_OK;
}

extern "C" rsp_status rsp_run_host(rsp_chain *c, const void *h_in, uint32_t n_chirps, uint32_t n_rx, void *h_out, uint32_t *h_peak_mask,
                                   rsp_detection *h_dets, uint32_t max_dets, uint32_t *h_n_dets) {
  if (!c || !h_in) return fail(RSP_EINVAL, "null argument");
  uint32_t rows;
  rsp_status s = rows_of(n_chirps, n_rx, &rows);
  if (s != RSP_OK) return s;
  RSP_ON_DEVICE(c);
  const bool fixed = c->p.arith == RSP_ARITH_FIXED;
  const int S = c->p.fft.runTime ? (int)c->r.fft[RSP_FFT_REG_FFTSIZE] : log2_ceil((uint32_t)c->p.fft.numPoints);
  if (S < 1 || S > 14) return fail(RSP_EINVAL, "fftSize register out of range");
  const size_t N = (size_t)1 << S;
  const size_t row_in = N * (fixed ? 4 : 8), row_out = h_out ? N * 4 : 0;
  const size_t row_mask = (!fixed && h_peak_mask) ? ((N + 31) / 32) * 4 : 0;
  // chunks of ~16 MiB of input (42.7 GB/s each way on PCIe Gen5 against 39.9 at 8 MiB; both directions at once peak at 49.8) keep three copies/kernels in flight
  size_t rows_per_chunk = std::max<size_t>(1, pipe_chunk_bytes((size_t)rows * row_in) / row_in);
  rows_per_chunk = std::min<size_t>(rows_per_chunk, std::max<uint32_t>(rows, 1));
  const bool in_pinned = is_pinned(h_in), out_pinned = !h_out || is_pinned(h_out);
  const bool mask_pinned = !row_mask || is_pinned(h_peak_mask);
  s = ensure_pipe(c, rows_per_chunk, row_in, row_out, row_mask, !in_pinned, !(out_pinned && mask_pinned));
  if (s != RSP_OK) return s;
  rows_per_chunk = c->pipe_rows_cap >= rows_per_chunk ? rows_per_chunk : c->pipe_rows_cap;
  rsp_dets dets = {nullptr, nullptr, 0};
  if (h_dets && max_dets) {
    if (c->det_cap < max_dets) {
      CUDA_TRY(cudaDeviceSynchronize());
      cudaFree(c->d_det_list); c->d_det_list = nullptr; c->det_cap = 0;
      CUDA_TRY(cudaMalloc((void **)&c->d_det_list, (size_t)max_dets * sizeof(rsp_detection)));
      c->det_cap = max_dets;
    }
    if (!c->d_det_count) CUDA_TRY(cudaMalloc((void **)&c->d_det_count, sizeof(uint32_t)));
    CUDA_TRY(cudaMemsetAsync(c->d_det_count, 0, sizeof(uint32_t), c->pipe_stream[0]));
    CUDA_TRY(cudaStreamSynchronize(c->pipe_stream[0]));
    dets.list = c->d_det_list; dets.count = c->d_det_count; dets.capacity = max_dets;
  }
  if (!h_out && !dets.list && !row_mask) return fail(RSP_EINVAL, "no output requested");
  const unsigned char *src = static_cast<const unsigned char *>(h_in);
  unsigned char *dst = static_cast<unsigned char *>(h_out);
  unsigned char *mdst = reinterpret_cast<unsigned char *>(h_peak_mask);
  const int depth = pipe_depth();
  struct Pending { size_t row0 = 0, n = 0; bool live = false; } pend[N_PIPE];
  auto drain = [&](int slot) -> rsp_status {   // wait for the slot's chunk and unstage it
    if (!pend[slot].live) return RSP_OK;
    CUDA_TRY(cudaStreamSynchronize(c->pipe_stream[slot]));
    if (!out_pinned && dst) std::memcpy(dst + pend[slot].row0 * row_out, c->stage_out[slot], pend[slot].n * row_out);
    if (!mask_pinned && row_mask) std::memcpy(mdst + pend[slot].row0 * row_mask, c->stage_out[slot], pend[slot].n * row_mask);
    pend[slot].live = false;
    return RSP_OK;
  };
  size_t chunk = 0;
  // Chunk schedule: the first chunks ramp up (1/8, 1/4, 1/2 of a chunk) and the last ones ramp down the same way, so that the
  // time in which only one of the two PCIe directions is busy (first H2D, last D2H) is that of a small chunk, not a full one.
  const size_t ramp_min = std::max<size_t>(1, rows_per_chunk / 8);
  const bool ramp = pipe_ramp() && rows > 2 * rows_per_chunk;
  size_t n = 0;
  for (size_t row0 = 0; row0 < rows; row0 += n, ++chunk) {
    const int slot = (int)(chunk % (size_t)depth);
    const size_t remaining = rows - row0;
    n = std::min<size_t>(rows_per_chunk, remaining);
    if (ramp) {
      if (chunk < 3) n = std::min(n, std::max(ramp_min, rows_per_chunk >> (3 - chunk)));
      if (remaining <= 2 * rows_per_chunk && remaining > ramp_min) n = std::min(n, std::max(ramp_min, remaining / 2));
    }
    s = drain(slot);
    if (s != RSP_OK) return s;
    cudaStream_t st = c->pipe_stream[slot];
    const void *hsrc = src + row0 * row_in;
    if (!in_pinned) { std::memcpy(c->stage_in[slot], hsrc, n * row_in); hsrc = c->stage_in[slot]; }
    CUDA_TRY(cudaMemcpyAsync(c->pipe_in[slot], hsrc, n * row_in, cudaMemcpyHostToDevice, st));
    // detections carry the row index inside the call: the chunk's first row is passed as row_base
    s = run_stages(c, STAGE_FFT, STAGE_CFAR, c->pipe_in[slot], (uint32_t)n, row_out ? c->pipe_out[slot] : nullptr,
                   row_mask ? c->pipe_mask[slot] : nullptr, dets.list ? &dets : nullptr, st, (uint32_t)row0);
    if (s != RSP_OK) return s;
    if (row_out) {
      void *hdst = out_pinned ? (void *)(dst + row0 * row_out) : c->stage_out[slot];
      CUDA_TRY(cudaMemcpyAsync(hdst, c->pipe_out[slot], n * row_out, cudaMemcpyDeviceToHost, st));
    }
    if (row_mask) {
      if (!mask_pinned && !out_pinned && row_out) {   // one staging buffer: finish the threshold copy first
        CUDA_TRY(cudaStreamSynchronize(st));
        std::memcpy(dst + row0 * row_out, c->stage_out[slot], n * row_out);
      }
      void *hdst = mask_pinned ? (void *)(mdst + row0 * row_mask) : c->stage_out[slot];
      CUDA_TRY(cudaMemcpyAsync(hdst, c->pipe_mask[slot], n * row_mask, cudaMemcpyDeviceToHost, st));
      if (!mask_pinned && !out_pinned && row_out) {
        CUDA_TRY(cudaStreamSynchronize(st));
        std::memcpy(mdst + row0 * row_mask, c->stage_out[slot], n * row_mask);
        continue;   // already unstaged
      }
    }
    pend[slot].row0 = row0; pend[slot].n = n; pend[slot].live = true;
  }
  for (int i = 0; i < depth; ++i) { s = drain(i); if (s != RSP_OK) return s; }
  for (int i = 0; i < depth; ++i) CUDA_TRY(cudaStreamSynchronize(c->pipe_stream[i]));
  if (dets.list) {
    uint32_t cnt = 0;
    CUDA_TRY(cudaMemcpy(&cnt, c->d_det_count, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    const uint32_t ncopy = std::min(cnt, max_dets);
    if (ncopy) CUDA_TRY(cudaMemcpy(h_dets, c->d_det_list, (size_t)ncopy * sizeof(rsp_detection), cudaMemcpyDeviceToHost));
    if (h_n_dets) *h_n_dets = cnt;
  } else if (h_n_dets) *h_n_dets = 0;
  return RSP_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// introspection
extern "C" const char *rsp_status_str(rsp_status s) {
  switch (s) {
  case RSP_OK: return "RSP_OK";
  case RSP_EINVAL: return "RSP_EINVAL";
  case RSP_ECUDA: return "RSP_ECUDA";
  case RSP_ENOMEM: return "RSP_ENOMEM";
  case RSP_EUNSUPPORTED: return "RSP_EUNSUPPORTED";
  default: return "RSP_?";
  }
}
extern "C" const char *rsp_last_error(void) { return g_last_error.c_str(); }
extern "C" int rsp_abi_version(void) { return RSP_ABI_VERSION; }
extern "C" uint64_t rsp_kernel_launches(const rsp_chain *c) { return c ? c->launches : 0; }

extern "C" const char *rsp_variant_name(rsp_chain *c) {
  if (!c) return "";
  std::string err;
  Plan pl;
  char buf[96];
  if (c->p.arith == RSP_ARITH_FIXED) {
    FxArgs a;
    if (plan_fixed(c->p, c->r, c->t, STAGE_FFT, STAGE_CFAR, a, pl, err) != RSP_OK) { c->variant = "invalid: " + err; return c->variant.c_str(); }
    a.out = reinterpret_cast<uint32_t *>(c);   // any non-null output: the name is that of the full chain with output beats requested
    bind_fixed(c, a);
    const char *m = pl.mode == MODE_FAST ? "fast" : (pl.mode == MODE_FAST3 ? "fast3" : pl.mode == MODE_FASTC ? "fast-convergent" : pl.mode == MODE_FAST3C ? "fast3-convergent" : (a.radix2 ? (a.dit ? "radix2-dit" : "radix2-dif") : (a.dit ? "radix22-dit" : "generic")));
    if (fast_chain_supported(pl, a)) std::snprintf(buf, sizeof(buf), a.core == CORE_LIS ? "chain_fixed_fast_gos<%d>" : (a.core == CORE_OS ? "chain_fixed_fast_os<%d>" : "chain_fixed_fast<%d>"), pl.S);
    else std::snprintf(buf, sizeof(buf), "chain_fixed<%d,%s>", pl.S, m);
  } else {
    F32Args a;
    if (plan_f32(c->p, c->r, c->t, STAGE_FFT, STAGE_CFAR, a, pl, err) != RSP_OK) { c->variant = "invalid: " + err; return c->variant.c_str(); }
    a.out = reinterpret_cast<float *>(c);   // any non-null output: the name is that of the full chain with thresholds requested
    if (fast_f32_chain_supported(pl, a)) std::snprintf(buf, sizeof(buf), "chain_f32_fast<%d>", pl.S);
    else std::snprintf(buf, sizeof(buf), "chain_f32<%d>", pl.S);
  }
  c->variant = buf;
  return c->variant.c_str();
}

extern "C" uint32_t rsp_fft_in_beat_bytes(rsp_chain *c) {
  if (!c) return 0;
  if (c->p.arith != RSP_ARITH_FIXED) return 8;   // float2
  return fft_is_wide(c->p.fft) ? 8 : 4;
}
extern "C" uint32_t rsp_fft_out_beat_bytes(rsp_chain *c) {
  if (!c) return 0;
  if (c->p.arith != RSP_ARITH_FIXED) return 8;
  if (!fft_is_wide(c->p.fft)) return 4;
  WideArgs a; std::string err;
  if (plan_wide(c->p, c->r, c->t, a, err) != RSP_OK) return 0;
  return 8u * (uint32_t)a.out_words;
}
extern "C" rsp_status rsp_fft_overflow_flags(rsp_chain *c, uint32_t *mask, void *stream) {
  if (!c || !mask) return fail(RSP_EINVAL, "null argument");
  *mask = 0;
  if (c->d_ovf == nullptr) return RSP_OK;   // overflowReg = false: no status register elaborated, reads 0
  RSP_ON_DEVICE(c);
  CUDA_TRY(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  CUDA_TRY(cudaMemcpy(mask, c->d_ovf, sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return RSP_OK;
}

extern "C" uint64_t rsp_bytes_per_chirp(rsp_chain *c) {
  if (!c) return 0;
  const int S = c->p.fft.runTime ? (int)c->r.fft[RSP_FFT_REG_FFTSIZE] : log2_ceil((uint32_t)c->p.fft.numPoints);
  const uint64_t N = 1ull << S;
  if (c->p.arith == RSP_ARITH_FIXED) return N * 8;        // Cat(re16, im16) in, Cat(threshold, bin, peak) out
  return N * 8 + N * 4 + ((N + 31) / 32) * 4;             // float2 in, float threshold + 1 bit out
}
