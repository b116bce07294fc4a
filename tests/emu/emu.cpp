// emu.cpp -- CPU emulation harness of the run-time-parameter kernels (TEST INFRASTRUCTURE, never part of the product library).
//
// The per-thread phases of rsp_device.cuh / rsp_device_f32.cuh / rsp_device_gen.cuh are host-callable (RSP_HD); this file compiles
// them with g++ and runs them thread by thread and phase by phase in the order the CUDA kernels of rsp_kernels.cuh do (a barrier
// between phases = a loop over all threads of the row per phase), on argument blocks built by the product's own planner
// (rsp_host.hpp: validate_params, regs_reset, build_tables, plan_fixed / plan_f32 / plan_generator).  tests/test_emu_cpu.py compares
// the results with the oracle, so the index maps, the integer arithmetic and the host planning are checked on a box without a GPU.
// What it cannot cover: the CUDA launch itself and the specialised kernels of rsp_fast*.cu / rsp_wide.cu (device-only code).
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../rsp_chains_b200/csrc/rsp_host.hpp"

using namespace rsp;

namespace {

template <int S, int MODE, int K>
void run_passes(const FxArgs &a, const uint32_t *gin, uint32_t *bufA, uint32_t *bufM) {
  for (int tl = 0; tl < Geo<S>::TPC; ++tl) fft_pass<S, MODE, K>(a, tl, gin, bufA, bufM);
  if constexpr (K + 1 < fx_n_passes(S)) run_passes<S, MODE, K + 1>(a, gin, bufA, bufM);
}

template <int S, int MODE>
void chain_fixed_rows(const FxArgs &a) {
  using G = Geo<S>;
  constexpr int N = G::N, TPC = G::TPC, PTS = G::PTS;
  std::vector<uint32_t> bufA(G::BUF), bufM(G::BUF);
  uint32_t *bufP = bufA.data();   // the prefix sums reuse the exchange buffer, as in chain_fixed_kernel
  for (uint32_t row = 0; row < a.n_rows; ++row) {
    const uint32_t *gin = a.in + (size_t)row * N;
    if (a.first_stage == STAGE_FFT) run_passes<S, MODE, 0>(a, gin, bufA.data(), bufM.data());
    else for (int n = 0; n < N; ++n) bufM[pad(n)] = gin[n];
    if (a.last_stage != STAGE_CFAR) {
      for (int n = 0; n < N; ++n) a.out[(size_t)row * N + n] = bufM[pad(n)];
      continue;
    }
    const bool need_sums = !((a.core == CORE_LIS && a.cfar_algorithm == 1) || a.core == CORE_OS);
    if (need_sums) {
      int offset = 0;
      std::vector<int> offs(TPC);
      for (int tl = 0; tl < TPC; ++tl) { offs[tl] = offset; offset += cfar_thread_total<S>(a, tl, bufM.data()); }
      for (int tl = 0; tl < TPC; ++tl) cfar_write_prefix<S>(a, tl, offs[tl], bufM.data(), bufP);
    }
    for (int tl = 0; tl < TPC; ++tl)
      for (int j = 0; j < PTS; ++j) cfar_bin<S>(a, row, tl + j * TPC, bufM.data(), bufP, a.out);
  }
}

template <int MODE>
bool dispatch_fixed(int S, const FxArgs &a) {
  switch (S) {
#define EMU_CASE(s) case s: chain_fixed_rows<s, MODE>(a); return true;
    EMU_CASE(1) EMU_CASE(2) EMU_CASE(3) EMU_CASE(4) EMU_CASE(5) EMU_CASE(6) EMU_CASE(7) EMU_CASE(8) EMU_CASE(9) EMU_CASE(10)
#undef EMU_CASE
  default: return false;
  }
}

template <int S, int K>
void run_passes_f32(const F32Args &a, const F2 *gin, F2 *bufA, float *bufM, F2 *gout_c) {
  for (int tl = 0; tl < GeoF<S>::TPC; ++tl) fft_pass_f32<S, K>(a, tl, gin, bufA, bufM, gout_c);
  if constexpr (K + 1 < fx_n_passes(S)) run_passes_f32<S, K + 1>(a, gin, bufA, bufM, gout_c);
}

template <int S>
void chain_f32_rows(const F32Args &a) {
  using G = GeoF<S>;
  constexpr int N = G::N, TPC = G::TPC, PTS = G::PTS;
  std::vector<F2> bufA(G::BUFA);
  std::vector<float> bufM(G::BUFM), bufT(G::BUFM);
  for (uint32_t row = 0; row < a.n_rows; ++row) {
    if (a.first_stage == STAGE_FFT) {
      run_passes_f32<S, 0>(a, a.in + (size_t)row * N, bufA.data(), bufM.data(), a.out_c ? a.out_c + (size_t)row * N : nullptr);
      if (a.last_stage == STAGE_FFT) continue;
    } else for (int n = 0; n < N; ++n) bufM[pad(n)] = a.in_mag[(size_t)row * N + n];
    if (a.last_stage != STAGE_CFAR) { for (int n = 0; n < N; ++n) a.out[(size_t)row * N + n] = bufM[pad(n)]; continue; }
    for (int tl = 0; tl < TPC; ++tl) {
      const uint32_t bits = cfar_thread_f32<S>(a, row, tl, bufM.data(), bufT.data());
      if (a.peak_mask)
        for (int j = 0; j < PTS; ++j) {
          const int n = tl * PTS + j;
          if ((bits >> j) & 1u) a.peak_mask[(size_t)row * G::MASKW + (n >> 5)] |= 1u << (n & 31);
        }
    }
    if (a.out) for (int n = 0; n < N; ++n) a.out[(size_t)row * N + n] = bufT[pad(n)];
  }
}

struct Handle {
  rsp_chain_params p;
  Regs r;
  Tables t;
};

int fill(Handle &h, const rsp_chain_params *p, const uint32_t *fft_regs, const uint32_t *mag_regs, const uint32_t *cfar_regs, float scaler_f32,
         std::string &err) {
  h.p = *p;
  const rsp_status vs = validate_params(h.p, err);
  if (vs != RSP_OK) return (int)vs;
  regs_reset(h.p, h.r);
  if (fft_regs) { h.r.fft[RSP_FFT_REG_FFTSIZE] = fft_regs[0]; h.r.fft[RSP_FFT_REG_FFTDIR] = fft_regs[1]; h.r.fft[RSP_FFT_REG_KEEPMSBORLSB] = fft_regs[2]; h.r.fft_dowin = fft_regs[3]; }
  if (mag_regs) h.r.mag[0] = mag_regs[0];
  if (cfar_regs) for (int i = 0; i < 12; ++i) h.r.cfar[i] = cfar_regs[i];
  h.r.scaler_f32 = scaler_f32;
  build_tables(h.p, h.t);
  return 0;
}

void set_err(char *dst, int n, const std::string &s) { if (dst && n > 0) { std::snprintf(dst, (size_t)n, "%s", s.c_str()); } }

}  // namespace

extern "C" {

// fixed-point stages first..last (STAGE_*) over n_rows rows.  fft_regs = {fftSize, fftDir, keepMSBorLSB, doWindowing}; cfar_regs = the 12
// registers of the CFAR block by index; win_ram (optional) = numPoints raw protoWin words (FFTBlockWithWindowing); ovf (optional)
// receives the overflow status word.  Returns an rsp_status (0 = ok, -1 = size / mode not instantiated in the harness).
int rsp_emu_run_fixed(const rsp_chain_params *p, const uint32_t *fft_regs, const uint32_t *mag_regs, const uint32_t *cfar_regs, int first, int last,
                      const uint32_t *in, uint32_t n_rows, uint32_t *out, rsp_detection *dets, uint32_t *det_count, uint32_t det_cap,
                      const int32_t *win_ram, uint32_t *ovf, int *mode_out, char *errbuf, int errlen) {
  Handle h;
  std::string err;
  int rc = fill(h, p, fft_regs, mag_regs, cfar_regs, 0.f, err);
  if (rc) { set_err(errbuf, errlen, err); return rc; }
  if (win_ram) h.t.win_ram.assign(win_ram, win_ram + h.p.fft.numPoints);
  build_window(h.p, h.t);
  FxArgs a;
  Plan pl;
  const rsp_status s = plan_fixed(h.p, h.r, h.t, first, last, a, pl, err);
  if (s != RSP_OK) { set_err(errbuf, errlen, err); return (int)s; }
  a.tw = h.t.tw.data(); a.tw2 = h.t.tw2.data(); a.log_tab = h.t.log_tab.data();
  a.win = h.t.win.empty() ? nullptr : h.t.win.data(); a.win4 = h.t.win4.empty() ? nullptr : h.t.win4.data();
  a.in = in; a.out = out; a.n_rows = n_rows; a.row_base = 0;
  a.det_list = reinterpret_cast<Detection *>(dets); a.det_count = det_count; a.det_cap = det_cap;
  uint32_t ovf_word = 0;
  a.ovf_flags = (h.p.fft.overflowReg && first == STAGE_FFT) ? &ovf_word : nullptr;
  if (mode_out) *mode_out = pl.mode;
  bool ok;
  switch (pl.mode) {
  case MODE_FAST: ok = dispatch_fixed<MODE_FAST>(pl.S, a); break;
  case MODE_FAST3: ok = dispatch_fixed<MODE_FAST3>(pl.S, a); break;
  case MODE_FASTC: ok = dispatch_fixed<MODE_FASTC>(pl.S, a); break;
  case MODE_FAST3C: ok = dispatch_fixed<MODE_FAST3C>(pl.S, a); break;
  case MODE_RADIX2: ok = dispatch_fixed<MODE_RADIX2>(pl.S, a); break;
  default: ok = dispatch_fixed<MODE_GENERIC>(pl.S, a); break;
  }
  if (!ok) { set_err(errbuf, errlen, "size not instantiated in the emulation harness"); return -1; }
  if (ovf) *ovf = ovf_word;
  return 0;
}

int rsp_emu_run_f32(const rsp_chain_params *p, const uint32_t *fft_regs, const uint32_t *mag_regs, const uint32_t *cfar_regs, float scaler, int first,
                    int last, const float *in, uint32_t n_rows, float *out, uint32_t *peak_mask, char *errbuf, int errlen) {
  Handle h;
  std::string err;
  int rc = fill(h, p, fft_regs, mag_regs, cfar_regs, scaler, err);
  if (rc) { set_err(errbuf, errlen, err); return rc; }
  build_window(h.p, h.t);
  F32Args a;
  Plan pl;
  const rsp_status s = plan_f32(h.p, h.r, h.t, first, last, a, pl, err);
  if (s != RSP_OK) { set_err(errbuf, errlen, err); return (int)s; }
  a.tw = h.t.twf.data(); a.win = h.t.winf.empty() ? nullptr : h.t.winf.data();
  if (first == STAGE_CFAR) a.in_mag = in; else a.in = reinterpret_cast<const F2 *>(in);
  if (last == STAGE_FFT) a.out_c = reinterpret_cast<F2 *>(out); else a.out = out;
  a.peak_mask = peak_mask; a.n_rows = n_rows; a.row_base = 0;
  switch (pl.S) {
  case 4: chain_f32_rows<4>(a); break;
  case 6: chain_f32_rows<6>(a); break;
  case 8: chain_f32_rows<8>(a); break;
  case 10: chain_f32_rows<10>(a); break;
  default: set_err(errbuf, errlen, "size not instantiated in the emulation harness"); return -1;
  }
  return 0;
}

// PLFG -> NCO: the planner's segment list evaluated sample by sample with the device code (closed form, running form, and the
// per-period phase table of the fused kernel).  plfg_regs = the whole PLFG register file, ram = the config RAM, nco_regs = {enable, factor}.
int rsp_emu_generate(const rsp_chain_params *p, const uint32_t *plfg_regs, int n_plfg_regs, const uint32_t *ram, int n_ram, const uint32_t *nco_regs,
                     uint64_t n_samples, uint32_t *nco_closed, uint32_t *nco_running, uint32_t *nco_table, int32_t *plfg_out, char *errbuf, int errlen) {
  Handle h;
  std::string err;
  int rc = fill(h, p, nullptr, nullptr, nullptr, 0.f, err);
  if (rc) { set_err(errbuf, errlen, err); return rc; }
  for (int i = 0; i < n_plfg_regs && i < (int)h.r.plfg.size(); ++i) h.r.plfg[(size_t)i] = plfg_regs[i];
  for (int i = 0; i < n_ram && i < (int)h.r.plfg_ram.size(); ++i) h.r.plfg_ram[(size_t)i] = ram[i];
  if (nco_regs) { h.r.nco[0] = nco_regs[0]; h.r.nco[1] = nco_regs[1]; }
  GenPlan gp;
  const rsp_status s = plan_generator(h.p, h.r, gp, err);
  if (s != RSP_OK) { set_err(errbuf, errlen, err); return (int)s; }
  if (n_samples > gp.total) { set_err(errbuf, errlen, "the PLFG schedule ends before the requested samples"); return RSP_EINVAL; }
  GenArgs g;
  std::memset(&g, 0, sizeof(g));
  g.segs = gp.segs.data(); g.n_segs = (uint32_t)gp.segs.size(); g.period_len = gp.period_len; g.period_phase = gp.period_phase;
  g.nco_tab = h.t.nco_tab.data(); g.nco_full = h.t.nco_full.empty() ? nullptr : h.t.nco_full.data();
  g.table_size = h.p.nco.tableSize; g.addr_bits = log2_ceil((uint32_t)h.p.nco.tableSize); g.table_width = h.p.nco.tableWidth;
  g.phase_width = h.p.nco.phaseWidth; g.out_width = h.p.plfg.outputWidthInt + h.p.plfg.outputWidthFrac;
  g.phase_acc_enable = h.p.nco.phaseAccEnable ? 1 : 0;
  g.rasterized = h.p.nco.rasterizedMode ? 1 : 0;
  g.mult_trim = h.p.nco.roundingMode;
  if (h.p.nco.useMultiplier) {
    const int w = h.p.nco.tableWidth;
    g.mult_f = (int)wrap_host((h.r.nco[0] & 1u) ? (long long)h.r.nco[1] : (1LL << 14), w);
    g.mult_on = (g.mult_f != (1 << (w - 2))) ? 1 : 0;
  }
  GenState st;
  for (uint64_t k = 0; k < n_samples; ++k) {
    int v = 0;
    if (nco_closed) nco_closed[k] = gen_sample(g, k, &v);
    if (plfg_out) { if (!nco_closed) gen_sample(g, k, &v); plfg_out[k] = v; }
    if (nco_running) {
      if (k % 8 == 0) gen_start(g, k, st); else gen_next(g, k, st);   // runs of 8 as generator_kernel does
      int v2;
      nco_running[k] = gen_beat(g, st, &v2);
    }
    if (nco_table) {   // the fused kernel's form: phase table of one period + beat table (0 when the configuration has none)
      if (gp.phase_tab.empty() || h.t.nco_full.empty()) nco_table[k] = 0;
      else {
        const uint64_t per = k / gp.period_len, q = k - per * gp.period_len;
        const uint32_t mask = (1u << h.p.nco.phaseWidth) - 1u;
        nco_table[k] = h.t.nco_full[((uint32_t)gp.phase_tab[(size_t)q] + (uint32_t)per * gp.period_phase) & mask];
      }
    }
  }
  return 0;
}

}  // extern "C"
