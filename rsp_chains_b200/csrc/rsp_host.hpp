// rsp_host.hpp -- host-side state of a chain: parameter checks (the reference's `require`s), register files with the
// reference's field widths and reset values, table generation, and the translation of (params, registers) into the
// kernel argument blocks.  Included by the translation units of the C-ABI library (rsp_abi.cu, rsp_fast*.cu).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rsp_chains.h"
#include "rsp_device.cuh"
#include "rsp_device_f32.cuh"
#include "rsp_device_gen.cuh"
#include "rsp_wide.cuh"

namespace rsp {

inline bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }
inline int popcount32(uint32_t v) { int n = 0; while (v) { n += (int)(v & 1u); v >>= 1; } return n; }
// the FFT block needs beats wider than Cat(re16, im16): the 64-bit kernel of rsp_wide.cu (FFT block alone)
inline bool fft_is_wide(const rsp_fft_params &f) { return f.dataWidth > 16 || f.twiddleWidth > 16 || f.expandLogic != 0; }
inline int log2_ceil(uint32_t v) { int n = 0; while ((1u << n) < v) ++n; return n; }   // chisel3.util.log2Ceil
inline int bit_length(uint32_t v) { int n = 0; while (v) { ++n; v >>= 1; } return n; }
inline long long wrap_host(long long v, int w) { return wrapw64(v, w); }
inline long long from_double(double x, int bp) { return (long long)std::floor(std::ldexp(x, bp) + 0.5); } // FixedPoint.fromDouble

struct Regs {
  uint32_t nco[2] = {0, 0};   // enableMultiplying, multiplyingFactor (useMultiplier; NCOLazyModuleBlock.scala:242-243)
  uint32_t fft_dowin = 0;   // doWindowing of FFTBlockWithWindowing (FFTBlockWithWindowing.scala:109,141)
  uint32_t fft[5];
  uint32_t mag[1];
  uint32_t cfar[12];
  std::vector<uint32_t> plfg;      // 6 + maxChirpOrdinalNum + 2 * maxNumOfDifferentChirps words
  std::vector<uint32_t> plfg_ram;  // maxChirpOrdinalNum * maxNumOfSegments words
  float scaler_f32;
};

struct Tables {
  std::vector<uint32_t> tw;   // fixed twiddles, packed (cos16 << 16 | sin16 & 0xFFFF); per stageN three runs [W^2k | W^k | W^3k]
  std::vector<F2> twf;        // fp32 twiddles, same layout
  std::vector<float> twf4;    // fp32 twiddles again as (c, s, -s, c) quads: the two operands of the packed complex product (rsp_fast_f32.cu)
  std::vector<int32_t> tww;   // fixed twiddles as raw int32 pairs (c, s) in Q2.(twiddleWidth - 2), any twiddleWidth up to 32 (rsp_wide.cu)
  std::vector<int32_t> tw2;   // fixed twiddles again as int32 pairs (c, s) << (16 - (twiddleWidth - 2)): the fast kernels'
                              // product then carries its rounded, wrapped result in bits 16..31 (rsp_fast.cu)
  int tw_off[16];
  std::vector<int32_t> log_tab;
  std::vector<int32_t> nco_tab;
  std::vector<uint32_t> nco_full;   // beat of every phase word (GenArgs::nco_full), empty when phaseWidth > 12
  std::vector<int32_t> win;   // window ROM, protoWin raw values (empty: WindowFunctionTypes.None)
  std::vector<int32_t> win4;  // win << (16 - (windowWidth - 2)) for the fast fixed kernels
  std::vector<float> winf;    // the same coefficients as floats (fp32 flavour: protoWin = DspReal)
  std::vector<double> win_user;   // WindowFunctionTypes.User coefficients (rsp_fft_window_user)
  std::vector<int32_t> win_ram;   // FFTBlockWithWindowing: the coefficient RAM, raw protoWin values (rsp_fft_window_ram_write)
};

// ---- the reference's elaboration-time checks -------------------------------------------------------------------
inline rsp_status validate_params(const rsp_chain_params &p, std::string &err) {
  const rsp_fft_params &f = p.fft;
  const rsp_cfar_params &c = p.cfar;
  const rsp_mag_params &m = p.mag;
#define RSP_REQ(cond, st, msg) do { if (!(cond)) { err = msg; return st; } } while (0)
  RSP_REQ(p.arith == RSP_ARITH_FIXED || p.arith == RSP_ARITH_FP32, RSP_EINVAL, "arith must be RSP_ARITH_FIXED or RSP_ARITH_FP32");
  // FFTParams (SDFFFTParams.scala:53-86)
  RSP_REQ(is_pow2(f.numPoints), RSP_EINVAL, "number of points must be a power of 2");
  RSP_REQ(f.numPoints >= 2 && f.numPoints <= 16384, RSP_EUNSUPPORTED, "numPoints must be in 2..16384 (one chirp stays resident in shared memory)");
  RSP_REQ(f.sdfRadix == 2 || f.sdfRadix == 4, RSP_EINVAL, "Radix must be 2 or 2^2");   // SDFFFTParams.scala:61-64
  RSP_REQ(f.decimType == 0 || f.decimType == 1, RSP_EINVAL, "decimType must be DIF or DIT");
  RSP_REQ(!f.expandLogic || (f.sdfRadix == 4 && f.decimType == 0 && p.arith == RSP_ARITH_FIXED), RSP_EUNSUPPORTED,
          "expandLogic (grow) stages are built for the fixed-point radix 2^2 DIF chain only");
  RSP_REQ(f.windowFunc >= RSP_WINDOW_NONE && f.windowFunc <= RSP_WINDOW_USER, RSP_EINVAL, "unknown windowFunc");
  RSP_REQ(!f.windowRam || f.windowFunc == RSP_WINDOW_NONE, RSP_EINVAL, "Parameter for window function inside SDFFFT module must be set to None type");   // FFTBlockWithWindowing.scala:19
  RSP_REQ(!f.windowRam || p.arith == RSP_ARITH_FIXED, RSP_EUNSUPPORTED, "FFTBlockWithWindowing is built for the fixed-point flavour");
  RSP_REQ(!f.windowRam || (f.windowWidth >= 3 && f.windowWidth <= 16), RSP_EUNSUPPORTED, "window dataWidth must be in 3..16");
  RSP_REQ(f.windowFunc == RSP_WINDOW_NONE || (f.windowWidth >= 3 && f.windowWidth <= 16), RSP_EUNSUPPORTED, "window dataWidth must be in 3..16");
  if (p.arith == RSP_ARITH_FIXED) {
    RSP_REQ(f.dataWidth >= 2 && f.dataWidth <= 32, RSP_EUNSUPPORTED, "dataWidth must be in 2..32");
    RSP_REQ(f.binPoint >= 0 && f.binPoint < f.dataWidth, RSP_EINVAL, "binPoint must be in 0..dataWidth-1");
    RSP_REQ(f.twiddleWidth >= 3 && f.twiddleWidth <= 32, RSP_EUNSUPPORTED, "twiddleWidth must be in 3..32");
    if (fft_is_wide(f)) {   // wide beats: the FFT block alone (rsp_fft_run) on the 64-bit kernel
      RSP_REQ(f.sdfRadix == 4 && f.decimType == 0, RSP_EUNSUPPORTED, "wide data / twiddles / grow stages are built for the radix 2^2 DIF chain only");
      RSP_REQ(f.numPoints <= 8192, RSP_EUNSUPPORTED, "the wide FFT block keeps a row of 64-bit points in shared memory: numPoints <= 8192");
      RSP_REQ(f.dataWidth + popcount32(f.expandLogic) <= 40, RSP_EUNSUPPORTED, "stage types wider than 40 bits are not built");
      RSP_REQ(f.windowFunc == RSP_WINDOW_NONE || f.windowWidth <= 16, RSP_EUNSUPPORTED, "window dataWidth must be in 3..16");
    }
    RSP_REQ(f.trimType >= 0 && f.trimType <= RSP_TRIM_NO_TRIM, RSP_EINVAL, "unknown trimType");
    RSP_REQ(m.dataWidth >= 2 && m.dataWidth <= 16 && m.binPoint >= 0 && m.binPoint <= 15, RSP_EUNSUPPORTED, "MAGParams proto must fit 16 bits");
    RSP_REQ(m.dataWidthLog >= 2 && m.dataWidthLog <= 32 && m.binPointLog >= 0 && m.binPointLog < 31, RSP_EINVAL, "bad protoLog");
    RSP_REQ(m.log2LookUpWidth >= 1 && m.log2LookUpWidth <= 14 && m.log2LookUpWidth <= m.dataWidth - 1, RSP_EINVAL, "log2LookUpWidth must be in 1..dataWidth-1");
    RSP_REQ(c.inWidth >= 2 && c.inWidth <= 16 && c.inBinPoint >= 0 && c.inBinPoint <= 15, RSP_EUNSUPPORTED, "CFAR protoIn must fit 16 bits");
    RSP_REQ(c.thresholdWidth >= 2 && c.thresholdWidth <= 32 && c.thresholdBinPoint >= 0 && c.thresholdBinPoint <= 31, RSP_EINVAL, "bad protoThreshold");
    RSP_REQ(c.scalerWidth >= 2 && c.scalerWidth <= 16 && c.scalerBinPoint >= 0 && c.scalerBinPoint <= 15, RSP_EUNSUPPORTED, "CFAR protoScaler must fit 16 bits");
  }
  // CFARParams (CFARParams.scala:40-44)
  RSP_REQ(is_pow2(c.leadLaggWindowSize) && is_pow2(c.fftSize), RSP_EINVAL, "leadLaggWindowSize and fftSize must be powers of 2");
  RSP_REQ(c.leadLaggWindowSize > 0 && c.guardWindowSize > 0, RSP_EINVAL, "leadLaggWindowSize and guardWindowSize must be positive");
  RSP_REQ(c.leadLaggWindowSize > c.guardWindowSize, RSP_EINVAL, "leadLaggWindowSize must exceed guardWindowSize");
  RSP_REQ(c.fftSize <= 16384 && c.leadLaggWindowSize <= 4096, RSP_EUNSUPPORTED, "CFAR fftSize must be <= 16384");
  RSP_REQ(c.CFARAlgorithm >= RSP_CFAR_CA && c.CFARAlgorithm <= RSP_CFAR_OS_GOLDEN, RSP_EINVAL, "unknown CFARAlgorithm");
  if (c.includeCASH)
    RSP_REQ(c.minSubWindowSize > 0, RSP_EINVAL, "If includeCash parameter is set then minSubWindowSize needs to be defined as a positive value");
  if (p.hasGenerator) {
    const rsp_plfg_params &g = p.plfg;
    const rsp_nco_params &n = p.nco;
    // FixedPLFGParams (PLFGDspBlockMem.scala:41-48)
    RSP_REQ(g.outputWidthInt >= 2 && g.outputWidthFrac >= 0, RSP_EINVAL, "outputWidthInt >= 2 and outputWidthFrac >= 0 required");
    RSP_REQ(is_pow2(g.maxNumOfSegments) && is_pow2(g.maxNumOfDifferentChirps) && is_pow2(g.maxNumOfRepeatedChirps) &&
            is_pow2(g.maxChirpOrdinalNum) && is_pow2(g.maxNumOfFrames), RSP_EINVAL, "PLFG maxima must be powers of 2");
    RSP_REQ(g.outputWidthInt + g.outputWidthFrac <= 16 && g.maxNumOfSamplesWidth >= 1 &&
            g.outputWidthInt + g.maxNumOfSamplesWidth + 2 <= 32, RSP_EUNSUPPORTED, "PLFG config word must fit 32 bits");
    RSP_REQ(g.maxNumOfSegments <= 64 && g.maxNumOfDifferentChirps <= 256 && g.maxChirpOrdinalNum <= 256, RSP_EUNSUPPORTED, "PLFG tables too large");
    // FixedNCOParams (NCO.scala:54-55)
    RSP_REQ(n.tableWidth >= 3 && n.tableWidth <= 16, RSP_EINVAL, "tableWidth must be in 3..16");
    RSP_REQ(n.tableSize >= 1 && n.tableSize <= 65536, RSP_EINVAL, "tableSize must be in 1..65536");
    if (n.rasterizedMode) {   // NCOTableRasterizedMode: any table size; the phase width is not a parameter (NCO.scala:56-62, NCOTable.scala:48-54)
      RSP_REQ(n.phaseWidth == log2_ceil((uint32_t)n.tableSize) + 2, RSP_EINVAL, "in rasterized mode phaseWidth = log2Ceil(tableSize) + 2 (NCO.scala:57-58)");
    } else {
      RSP_REQ(is_pow2(n.tableSize), RSP_EINVAL, "Table Size must be a power of 2 when working in standard mode");   // NCOTable.scala:103
      RSP_REQ(n.phaseWidth >= log2_ceil((uint32_t)n.tableSize) + 2 && n.phaseWidth <= 32, RSP_EINVAL, "phaseWidth must cover quadrant + table address");
    }
    RSP_REQ(!n.nInterpolationTerms && !n.ditherEnable && !n.useQAM, RSP_EUNSUPPORTED, "NCO Taylor terms, dither and QAM are not built");
    RSP_REQ(n.roundingMode >= 0 && n.roundingMode <= RSP_TRIM_NO_TRIM, RSP_EINVAL, "unknown NCO roundingMode");
    RSP_REQ(p.arith == RSP_ARITH_FIXED, RSP_EUNSUPPORTED, "the PLFG/NCO generator feeds the fixed-point chain only");
  }
#undef RSP_REQ
  return RSP_OK;
}

// ---- register files ---------------------------------------------------------------------------------------------
inline int plfg_reg_count(const rsp_plfg_params &g) { return 6 + g.maxChirpOrdinalNum + 2 * g.maxNumOfDifferentChirps; }

inline void regs_reset(const rsp_chain_params &p, Regs &r) {
  const int numStages = log2_ceil((uint32_t)p.fft.numPoints);
  r.fft_dowin = 0; r.nco[0] = r.nco[1] = 0;
  std::memset(r.fft, 0, sizeof(r.fft)); std::memset(r.mag, 0, sizeof(r.mag)); std::memset(r.cfar, 0, sizeof(r.cfar));
  r.fft[RSP_FFT_REG_FFTSIZE] = (uint32_t)numStages;           // FFTBlock.scala:62
  r.fft[RSP_FFT_REG_FFTDIR] = 1;                               // :63
  r.fft[RSP_FFT_REG_KEEPMSBORLSB] = 0;                         // :64
  r.cfar[RSP_CFAR_REG_FFTWIN] = (uint32_t)p.cfar.fftSize;      // CFARDspBlock.scala:29-43
  r.cfar[RSP_CFAR_REG_WINDOWCELLS] = (uint32_t)p.cfar.leadLaggWindowSize;
  r.cfar[RSP_CFAR_REG_GUARDCELLS] = (uint32_t)p.cfar.guardWindowSize;
  r.cfar[RSP_CFAR_REG_INDEXLEAD] = (uint32_t)p.cfar.leadLaggWindowSize / 2;
  r.cfar[RSP_CFAR_REG_INDEXLAGG] = (uint32_t)p.cfar.leadLaggWindowSize / 2;
  r.cfar[RSP_CFAR_REG_SUBWINDOWSIZE] = (uint32_t)p.cfar.leadLaggWindowSize;
  r.scaler_f32 = 0.0f;
  r.plfg.clear(); r.plfg_ram.clear();
  if (p.hasGenerator) {
    r.plfg.assign((size_t)plfg_reg_count(p.plfg), 0u);
    r.plfg[RSP_PLFG_REG_FRAMENUM] = 1;                         // PLFGDspBlockMem.scala:133-137
    r.plfg[RSP_PLFG_REG_INTERFRAMENUMOFSAMPLES] = 63u & ((1u << p.plfg.maxNumOfSamplesWidth) - 1);
    r.plfg[RSP_PLFG_REG_DIFFERENTCHIRPSNUM] = 1;
    for (int i = 0; i < p.plfg.maxChirpOrdinalNum; ++i)       // segmentNums reset to 1 (:139)
      r.plfg[RSP_PLFG_REG_SEGMENTNUMS0 + i] = 1u & ((1u << log2_ceil((uint32_t)p.plfg.maxNumOfSegments)) - 1);
    r.plfg_ram.assign((size_t)p.plfg.maxChirpOrdinalNum * p.plfg.maxNumOfSegments, 0u);
  }
}

// width in bits of a register field (RegField widths of the reference regmaps); -1: no such register; ro: read-only
inline int reg_width(const rsp_chain_params &p, rsp_block blk, uint32_t idx, bool *ro) {
  *ro = false;
  const int numStages = log2_ceil((uint32_t)p.fft.numPoints);
  const rsp_cfar_params &c = p.cfar;
  switch (blk) {
  case RSP_BLK_FFT:
    switch (idx) {
    case RSP_FFT_REG_FFTSIZE: return log2_ceil((uint32_t)numStages + 1);  // the register's own width (FFTBlock.scala:62)
    case RSP_FFT_REG_FFTDIR: return 1;
    case RSP_FFT_REG_KEEPMSBORLSB: return numStages;
    case RSP_FFT_REG_BUSY: *ro = true; return 1;
    case RSP_FFT_REG_OVERFLOW: *ro = true; return log2_ceil((uint32_t)numStages);   // RegInit(0.U(log2Ceil(numStages).W)) (FFTBlock.scala:67,97)
    default: return -1;
    }
  case RSP_BLK_MAG: return idx == RSP_MAG_REG_SEL ? 2 : -1;
  case RSP_BLK_CFAR:
    switch (idx) {
    case RSP_CFAR_REG_FFTWIN: return log2_ceil((uint32_t)c.fftSize + 1);
    case RSP_CFAR_REG_THRESHOLDSCALER: return c.scalerWidth;
    case RSP_CFAR_REG_LOGORLINEARMODE: return 1;
    case RSP_CFAR_REG_DIVSUM: return log2_ceil((uint32_t)log2_ceil((uint32_t)c.leadLaggWindowSize + 1));
    case RSP_CFAR_REG_PEAKGROUPING: return 1;
    case RSP_CFAR_REG_CFARALGORITHM: return 1;
    case RSP_CFAR_REG_CFARMODE: return 2;
    case RSP_CFAR_REG_WINDOWCELLS: case RSP_CFAR_REG_INDEXLAGG: case RSP_CFAR_REG_INDEXLEAD: case RSP_CFAR_REG_SUBWINDOWSIZE:
      return log2_ceil((uint32_t)c.leadLaggWindowSize + 1);
    case RSP_CFAR_REG_GUARDCELLS: return log2_ceil((uint32_t)c.guardWindowSize + 1);
    default: return -1;
    }
  case RSP_BLK_PLFG: {
    if (!p.hasGenerator) return -1;
    const rsp_plfg_params &g = p.plfg;
    const uint32_t a = RSP_PLFG_REG_SEGMENTNUMS0, b = a + g.maxChirpOrdinalNum, d = b + g.maxNumOfDifferentChirps, e = d + g.maxNumOfDifferentChirps;
    if (idx == RSP_PLFG_REG_ENABLE || idx == RSP_PLFG_REG_RESET) return 1;
    if (idx == RSP_PLFG_REG_FRAMENUM) return log2_ceil((uint32_t)g.maxNumOfFrames);
    if (idx == RSP_PLFG_REG_INTERFRAMENUMOFSAMPLES) return g.maxNumOfSamplesWidth < 10 ? g.maxNumOfSamplesWidth : 10;
    if (idx == RSP_PLFG_REG_DIFFERENTCHIRPSNUM) return log2_ceil((uint32_t)g.maxNumOfDifferentChirps);
    if (idx == RSP_PLFG_REG_STARTINGPOINT) return g.outputWidthInt;
    if (idx >= a && idx < b) return log2_ceil((uint32_t)g.maxNumOfSegments);
    if (idx >= b && idx < d) return log2_ceil((uint32_t)g.maxNumOfRepeatedChirps);
    if (idx >= d && idx < e) return log2_ceil((uint32_t)g.maxChirpOrdinalNum);
    return -1;
  }
  case RSP_BLK_NCO:   // chain configuration: pinc Streaming, poff Fixed: registers exist with the multiplier only (NCOLazyModuleBlock.scala:246-271)
    if (!p.hasGenerator || !p.nco.useMultiplier) return -1;
    if (idx == RSP_NCO_REG_ENABLEMULTIPLYING) return 1;
    if (idx == RSP_NCO_REG_MULTIPLYINGFACTOR) return p.nco.tableWidth;
    return -1;
  default: return -1;
  }
}

inline uint32_t *reg_slot(Regs &r, rsp_block blk, uint32_t idx) {
  switch (blk) {
  case RSP_BLK_FFT: return &r.fft[idx];
  case RSP_BLK_MAG: return &r.mag[idx];
  case RSP_BLK_CFAR: return &r.cfar[idx];
  case RSP_BLK_PLFG: return &r.plfg[idx];
  case RSP_BLK_NCO: return &r.nco[idx];
  default: return nullptr;
  }
}

// ---- tables ------------------------------------------------------------------------------------------------------
// twiddle(t) = (cos, -sin)(2 pi t / stageN) through fromDoubleWithFixedWidth into Q2.(tw-2) (SDFChainRadix22.scala:255-267).
// Stored per stageN as three runs of q = stageN/4 entries, [W^2k | W^k | W^3k], i.e. the quarters 1..3 of the reference's
// index pattern [0]*q ++ [2k] ++ [k] ++ [3k] (:255-259), so that neighbouring lanes read neighbouring words.
inline void build_tables(const rsp_chain_params &p, Tables &t) {
  const int Smax = log2_ceil((uint32_t)p.fft.numPoints);
  const int tww = p.fft.twiddleWidth, twbp = tww - 2;
  t.tw.clear(); t.twf.clear(); t.tw2.clear(); t.twf4.clear(); t.tww.clear();
  const int pre = 16 - twbp > 0 ? 16 - twbp : 0;
  for (int L = 0; L < 16; ++L) t.tw_off[L] = 0;
  const double PI = 3.14159265358979323846;  // scala.math.Pi
  // the radix-2 chain reads run 0 of the table of stageN = 4 * delay as its stage ROM, W_(2 delay)^k = W_stageN^2k bit for bit
  // (SDFChainRadix2.scala:300-306): its first stage needs one table more
  const int Lmax = p.fft.sdfRadix == 2 ? (Smax + 1 < 15 ? Smax + 1 : 15) : Smax;
  for (int L = 3; L <= Lmax; ++L) {
    const int stageN = 1 << L, q = stageN / 4;
    t.tw_off[L] = (int)t.tw.size();
    const int mult[3] = {2, 1, 3};
    for (int slot = 0; slot < 3; ++slot)
      for (int k = 0; k < q; ++k) {
        const int tt = mult[slot] * k;
        const double ang = 2.0 * PI * (double)tt / (double)stageN;
        const double c = std::cos(ang), s = -std::sin(ang);
        const long long re = wrap_host(from_double(c, twbp), tww), im = wrap_host(from_double(s, twbp), tww);
        t.tw.push_back(pack_iq((int)re, (int)im));
        t.tww.push_back((int32_t)re); t.tww.push_back((int32_t)im);
        const int32_t c4 = (int32_t)(re * (1LL << pre)), s4 = (int32_t)(im * (1LL << pre));   // (meaningful for twiddleWidth <= 16 only)
        t.tw2.push_back(c4); t.tw2.push_back(s4);
        F2 f; f.x = (float)c; f.y = (float)s;
        t.twf.push_back(f);
        t.twf4.push_back(f.x); t.twf4.push_back(f.y); t.twf4.push_back(-f.y); t.twf4.push_back(f.x);
      }
  }
  if (t.tw.empty()) { t.tw.push_back(0); for (int z = 0; z < 2; ++z) { t.tw2.push_back(0); t.tww.push_back(0); } F2 f; f.x = 1.f; f.y = 0.f; t.twf.push_back(f);
                      t.twf4.push_back(1.f); t.twf4.push_back(0.f); t.twf4.push_back(0.f); t.twf4.push_back(1.f); }
  // log2 LUT (LogMagMux.scala:120-131)
  const int Lw = p.mag.log2LookUpWidth, num = 1 << Lw;
  t.log_tab.resize((size_t)num);
  const double ln2 = std::log(2.0);
  for (int n = 0; n < num; ++n) {
    const double v = std::log((double)n / (double)num + 1.0) / ln2;
    t.log_tab[(size_t)n] = (int32_t)wrap_host(from_double(v, p.mag.binPointLog), p.mag.binPointLog + 1);
  }
  // NCO quarter-wave sine table (NCOTable.scala:254-258)
  t.nco_tab.clear();
  if (p.hasGenerator) {
    t.nco_tab.resize((size_t)p.nco.tableSize);
    for (int i = 0; i < p.nco.tableSize; ++i)
      t.nco_tab[(size_t)i] = (int32_t)from_double(std::sin(((double)i * 0.5 * PI) / (double)p.nco.tableSize), p.nco.tableWidth - 2);
  } else t.nco_tab.push_back(0);
  t.nco_full.clear();
  if (p.hasGenerator && p.nco.phaseWidth <= 12) {
    GenArgs g;
    std::memset(&g, 0, sizeof(g));
    g.nco_tab = t.nco_tab.data();   // host pointer: nco_lookup is the same code on both sides
    g.table_size = p.nco.tableSize; g.addr_bits = log2_ceil((uint32_t)p.nco.tableSize);
    g.table_width = p.nco.tableWidth; g.phase_width = p.nco.phaseWidth; g.rasterized = p.nco.rasterizedMode ? 1 : 0;
    t.nco_full.resize((size_t)1 << p.nco.phaseWidth);
    for (uint32_t ph = 0; ph < (1u << p.nco.phaseWidth); ++ph) t.nco_full[ph] = nco_lookup(g, ph);
  }
}

// window ROM (SDFFFT.scala:53-66): the periodic forms of WindowFunc.scala:41-65 evaluated in double, then
// fromDoubleWithFixedWidth into protoWin = FixedPoint(w, w - 2) (SDFFFTParams.scala:118)
inline void build_window(const rsp_chain_params &p, Tables &t) {
  const rsp_fft_params &f = p.fft;
  t.win.clear(); t.winf.clear(); t.win4.clear();
  if (f.windowRam) {   // FFTBlockWithWindowing: the ROM is a RAM the host writes; SyncReadMem content before that: zeros
    const int n = f.numPoints, wbp = f.windowWidth - 2;
    if (t.win_ram.size() != (size_t)n) t.win_ram.assign((size_t)n, 0);
    for (int k = 0; k < n; ++k) {
      t.win.push_back(t.win_ram[(size_t)k]);
      t.win4.push_back((int32_t)(t.win.back() * (1 << (16 - wbp))));
      t.winf.push_back((float)std::ldexp((double)t.win.back(), -wbp));
    }
    return;
  }
  if (f.windowFunc == RSP_WINDOW_NONE) return;
  const int n = f.numPoints, wbp = f.windowWidth - 2;
  const double PI = 3.14159265358979323846;
  for (int k = 0; k < n; ++k) {
    double v = 1.0;
    const double x = (double)k / (double)n;
    if (n > 1) switch (f.windowFunc) {
      case RSP_WINDOW_HAMMING: v = f.windowA0 - f.windowA1 * std::cos(2 * PI * x); break;
      case RSP_WINDOW_HANNING: v = 0.5 * (1 - std::cos(2 * PI * x)); break;
      case RSP_WINDOW_BLACKMAN: v = f.windowA0 - f.windowA1 * std::cos(2 * PI * x) + f.windowA2 * std::cos(4 * PI * x); break;
      case RSP_WINDOW_TRIANGULAR: v = 1 - std::fabs(((double)k - n / 2.0) / (n / 2.0)); break;
      default: v = (size_t)k < t.win_user.size() ? t.win_user[(size_t)k] : 1.0; break;
    }
    t.win.push_back((int32_t)wrap_host(from_double(v, wbp), f.windowWidth));
    t.win4.push_back((int32_t)(t.win.back() * (1 << (16 - wbp))));
    t.winf.push_back((float)v);
  }
}

// ---- (params, registers) -> kernel arguments ------------------------------------------------------------------------
struct Plan {
  int S;          // log2 of the transform / window size
  int mode;       // MODE_* of the fixed kernels
  int first_stage, last_stage;
};

// the part of the CFAR set-up common to both arithmetics; fills core/W/G/... of either argument block
template <typename A>
inline rsp_status plan_cfar_common(const rsp_chain_params &p, const Regs &r, int N, A &a, std::string &err) {
  const rsp_cfar_params &c = p.cfar;
  a.core = (c.CFARAlgorithm == RSP_CFAR_CA) ? (c.includeCASH ? CORE_ASR : CORE_MEM) : (c.CFARAlgorithm == RSP_CFAR_OS_GOLDEN ? CORE_OS : CORE_LIS);   // CFARcore.scala:62-64
  a.W = (int)r.cfar[RSP_CFAR_REG_WINDOWCELLS];
  a.G = (int)r.cfar[RSP_CFAR_REG_GUARDCELLS];
  a.div_sum = (int)r.cfar[RSP_CFAR_REG_DIVSUM];
  a.peak_grouping = (int)r.cfar[RSP_CFAR_REG_PEAKGROUPING];
  a.cfar_mode = (int)r.cfar[RSP_CFAR_REG_CFARMODE];
  a.log_or_linear = (int)r.cfar[RSP_CFAR_REG_LOGORLINEARMODE];
  a.cfar_algorithm = c.CFARAlgorithm == RSP_CFAR_GOS ? 1 : (c.CFARAlgorithm == RSP_CFAR_GOSCA ? (int)r.cfar[RSP_CFAR_REG_CFARALGORITHM] : 0);
  a.idx_lagg = (int)r.cfar[RSP_CFAR_REG_INDEXLAGG];
  a.idx_lead = (int)r.cfar[RSP_CFAR_REG_INDEXLEAD];
  a.sub_cells = (int)r.cfar[RSP_CFAR_REG_SUBWINDOWSIZE];
  a.sub_log2 = 0; while ((2 << a.sub_log2) <= a.sub_cells) ++a.sub_log2;
  a.guard_window_size = c.guardWindowSize;
  a.send_cut = c.sendCut;
  a.binw = log2_ceil((uint32_t)c.fftSize);
  if ((int)r.cfar[RSP_CFAR_REG_FFTWIN] != N) { err = "CFAR fftWin register must equal the run-time transform size"; return RSP_EINVAL; }
  if (N > c.fftSize) { err = "fftWin exceeds CFARParams.fftSize"; return RSP_EINVAL; }
  // run-time asserts of the cores (CFARCoreWithMem.scala:31-32)
  if (a.W < 1 || a.W > c.leadLaggWindowSize) { err = "windowCells must be in 1..leadLaggWindowSize"; return RSP_EINVAL; }
  if (a.G < 0 || a.G > c.guardWindowSize) { err = "guardCells must be in 0..guardWindowSize"; return RSP_EINVAL; }
  if (a.core == CORE_MEM && a.G < 1) { err = "guardCells must be positive"; return RSP_EINVAL; }
  if (!(N > 2 * a.W + 2 * a.G + 1)) { err = "fftWin must exceed 2*windowCells + 2*guardCells + 1"; return RSP_EINVAL; }
  if (a.core == CORE_LIS && a.cfar_algorithm == 1 &&
      (a.idx_lagg < 1 || a.idx_lagg > a.W || a.idx_lead < 1 || a.idx_lead > a.W)) { err = "indexLagg/indexLead must be in 1..windowCells"; return RSP_EINVAL; }
  if (a.core == CORE_ASR && (a.sub_cells < 1 || a.sub_cells > a.W || !is_pow2(a.sub_cells) || (a.W % a.sub_cells))) {
    err = "subWindowSize must be a power of 2 dividing windowCells"; return RSP_EINVAL; }
  // Peak grouping with guardCells = 0 on the Lis / ASR cores: the reference substitutes the ends of the reference windows for the
  // missing guard cells (CFARCoreWithLis.scala:191-192, CFARCoreWithASR.scala:326-327); that substitution is not built
  if (a.core == CORE_OS && a.peak_grouping) { err = "peakGrouping does not exist in the cfarOS golden"; return RSP_EINVAL; }
  if (a.core != CORE_MEM && a.peak_grouping && a.G == 0) { err = "peakGrouping with guardCells = 0 is not built for the Lis / ASR cores"; return RSP_EUNSUPPORTED; }
  if (a.core != CORE_ASR && a.cfar_mode == 3) { /* MuxLookup default: smallestOf (CFARCoreWithMem.scala:138-141) */ }
  return RSP_OK;
}

inline rsp_status plan_fixed(const rsp_chain_params &p, const Regs &r, const Tables &t, int first_stage, int last_stage,
                             FxArgs &a, Plan &pl, std::string &err) {
  std::memset(&a, 0, sizeof(a));
  const rsp_fft_params &f = p.fft;
  const int Smax = log2_ceil((uint32_t)f.numPoints);
  int S;
  if (first_stage == STAGE_CFAR) {
    const int N = (int)r.cfar[RSP_CFAR_REG_FFTWIN];
    if (!is_pow2(N) || N < 2) { err = "CFAR fftWin must be a power of 2 here"; return RSP_EUNSUPPORTED; }
    S = log2_ceil((uint32_t)N);
  } else {
    S = f.runTime ? (int)r.fft[RSP_FFT_REG_FFTSIZE] : Smax;
    if (S < 1 || S > Smax) { err = "fftSize register must be in 1..log2(numPoints)"; return RSP_EINVAL; }
    if (f.sdfRadix == 4 && f.decimType == 0 && ((Smax - S) & 1)) { err = "radix 2^2 run-time size must drop an even number of stages (Radix22RunTimeSpec.scala:18-19)"; return RSP_EINVAL; }
  }
  if (first_stage == STAGE_FFT && fft_is_wide(f)) {
    err = "beats wider than Cat(re16, im16) (dataWidth / twiddleWidth > 16 or grow stages): only the FFT block alone runs (rsp_fft_run); LogMagMux and CFAR take 16-bit cells";
    return RSP_EUNSUPPORTED;
  }
  pl.S = S; pl.first_stage = first_stage; pl.last_stage = last_stage;
  a.first_stage = first_stage; a.last_stage = last_stage;
  a.data_width = f.dataWidth; a.bp = f.binPoint; a.tw_bp = f.twiddleWidth - 2; a.trim = f.trimType; a.use4muls = f.use4Muls ? 1 : 0;
  a.radix2 = f.sdfRadix == 2 ? 1 : 0;
  a.dit = f.decimType == 1 ? 1 : 0;
  // radix-2^2 DIT: isMulTw = (index + 1) even for an odd number of elaborated stages, odd for an even one (SDFChainRadix22.scala:229)
  a.dit_tw_par = (Smax & 1) ? 1 : 0;
  // DIT runs natural-in / bit-reversed-out like DIF (rsp_device.cuh): its result always lands at the bit-reversed index, and
  // without useBitReverse the caller's beats are in bit-reversed order (SDFFFT.scala:217-270)
  a.bitrev = (f.useBitReverse || a.dit) ? 1 : 0;
  a.load_rev = (a.dit && !f.useBitReverse) ? 1 : 0;
  a.fft_dir = f.fftDirReg ? (int)(r.fft[RSP_FFT_REG_FFTDIR] & 1u) : (f.fftDir ? 1 : 0);
  // IFFT output scalar (SDFChainRadix22.scala:362, SDFChainRadix2.scala:250): 1.0 unless a stage grows or keepMSBorLSBReg is elaborated;
  // then it is the literal fromDouble(1.0 / numPoints), whose binary point is dsptools-context dependent and pinned by nothing
  if (first_stage == STAGE_FFT && !a.fft_dir && f.keepMSBorLSBReg) {
    err = "IFFT with keepMSBorLSBReg: the 1/numPoints output scalar of the reference is not built"; return RSP_EUNSUPPORTED; }
  const uint32_t phys = f.keepMSBorLSBReg ? r.fft[RSP_FFT_REG_KEEPMSBORLSB] : f.keepMSBorLSB;
  const uint32_t smask = S >= 32 ? 0xFFFFFFFFu : ((1u << S) - 1u);
  // a run-time size uses the LAST S stages in DIF and the FIRST S in DIT (SDFChainRadix2.scala:127-134,196-203)
  a.keep_mask = (first_stage == STAGE_CFAR) ? smask : (((a.dit ? phys : (phys >> (Smax - S)))) & smask);
  for (int L = 0; L < 16; ++L) a.tw_off[L] = t.tw_off[L];
  // windowing happens on the forward transform only (SDFFFT.scala:229-245: `when (fftDir)`); win itself is bound by the caller
  a.win_on = (first_stage == STAGE_FFT && f.windowFunc != RSP_WINDOW_NONE && a.fft_dir == 1) ? 1 : 0;
  // FFTBlockWithWindowing multiplies in front of the core whenever doWindowing is set, in both directions (FFTBlockWithWindowing.scala:159-166)
  if (f.windowRam) a.win_on = (first_stage == STAGE_FFT && (r.fft_dowin & 1u)) ? 1 : 0;
  a.win_shift = f.windowWidth - 2;
  a.sel = (int)(r.mag[RSP_MAG_REG_SEL] & 3u);
  a.mag_w = p.mag.dataWidth; a.mag_bp = p.mag.binPoint; a.log_w = p.mag.dataWidthLog; a.log_bp = p.mag.binPointLog;
  a.log_L = p.mag.log2LookUpWidth; a.mag_trim = TRIM_FLOOR;   // DspContext default trimType
  const rsp_cfar_params &c = p.cfar;
  a.in_w = c.inWidth; a.in_bp = c.inBinPoint; a.thr_w = c.thresholdWidth; a.thr_bp = c.thresholdBinPoint;
  a.sc_w = c.scalerWidth; a.sc_bp = c.scalerBinPoint;
  a.sumw = c.inWidth + bit_length((uint32_t)log2_ceil((uint32_t)c.leadLaggWindowSize)) + 1;   // CFARCoreWithMem.scala:34
  a.scaler = (int)wrap_host((long long)r.cfar[RSP_CFAR_REG_THRESHOLDSCALER], c.scalerWidth);
  a.cfar_trim = TRIM_FLOOR;
  if (last_stage == STAGE_CFAR) {
    const rsp_status st = plan_cfar_common(p, r, 1 << S, a, err);
    if (st != RSP_OK) return st;
    if (a.sumw > 31) { err = "CFAR sum register too wide"; return RSP_EUNSUPPORTED; }
  }
  // kernel variant
  const int bp_a = a.bp, bp_b = a.tw_bp;
  const int s1 = bp_a + bp_b - ((bp_a > bp_b ? bp_a : bp_b) + 1);
  const bool shape = f.dataWidth == 16 && a.keep_mask == smask && a.fft_dir == 1 && p.mag.dataWidth == 16;
  const bool common = shape && f.trimType == RSP_TRIM_ROUND_HALF_UP;
  const bool common_c = shape && f.trimType == RSP_TRIM_CONVERGENT;   // (RSP_TRIM_CONVERGENT also stands for RoundHalfToEven)
  pl.mode = MODE_GENERIC;
  a.ovf_shift = a.dit ? 0 : Smax - S;    // physical stage of this transform's stage 0 (a run-time size uses the last S stages in DIF)
  if (a.radix2 || a.dit) { pl.mode = MODE_RADIX2; return RSP_OK; }   // run-time-parameter kernel with the radix-2 stage structure
  if (f.overflowReg && first_stage == STAGE_FFT) return RSP_OK;      // the overflow status is computed by the run-time-parameter kernels
  if (common && s1 <= 0) pl.mode = MODE_FAST;
  else if (common && s1 >= 1 && !f.use4Muls) pl.mode = MODE_FAST3;
  else if (common_c && s1 <= 0) pl.mode = MODE_FASTC;
  else if (common_c && s1 >= 1 && !f.use4Muls) pl.mode = MODE_FAST3C;
  return RSP_OK;
}

inline rsp_status plan_f32(const rsp_chain_params &p, const Regs &r, const Tables &t, int first_stage, int last_stage,
                           F32Args &a, Plan &pl, std::string &err) {
  std::memset(&a, 0, sizeof(a));
  const rsp_fft_params &f = p.fft;
  const int Smax = log2_ceil((uint32_t)f.numPoints);
  int S;
  if (first_stage == STAGE_CFAR) {
    const int N = (int)r.cfar[RSP_CFAR_REG_FFTWIN];
    if (!is_pow2(N) || N < 2) { err = "CFAR fftWin must be a power of 2 here"; return RSP_EUNSUPPORTED; }
    S = log2_ceil((uint32_t)N);
  } else {
    S = f.runTime ? (int)r.fft[RSP_FFT_REG_FFTSIZE] : Smax;
    if (S < 1 || S > Smax) { err = "fftSize register must be in 1..log2(numPoints)"; return RSP_EINVAL; }
  }
  pl.S = S; pl.mode = 0; pl.first_stage = first_stage; pl.last_stage = last_stage;
  a.first_stage = first_stage; a.last_stage = last_stage;
  // DIT (either radix): natural output always; bit-reversed input unless the ping-pong buffer is elaborated in front
  a.bitrev = (f.useBitReverse || f.decimType == 1) ? 1 : 0;
  a.load_rev = (f.decimType == 1 && !f.useBitReverse) ? 1 : 0;
  a.fft_dir = f.fftDirReg ? (int)(r.fft[RSP_FFT_REG_FFTDIR] & 1u) : (f.fftDir ? 1 : 0);
  for (int L = 0; L < 16; ++L) a.tw_off[L] = t.tw_off[L];
  a.win_on = (first_stage == STAGE_FFT && f.windowFunc != RSP_WINDOW_NONE && a.fft_dir == 1) ? 1 : 0;
  a.sel = (int)(r.mag[RSP_MAG_REG_SEL] & 3u);
  a.scaler = r.scaler_f32;
  if (last_stage == STAGE_CFAR) {
    const rsp_status st = plan_cfar_common(p, r, 1 << S, a, err);
    if (st != RSP_OK) return st;
    a.inv_div = std::ldexp(1.0f, -a.div_sum);
  }
  return RSP_OK;
}

// the FFT block alone with wide beats (rsp_wide.cu): radix 2^2 DIF, any dataWidth / twiddleWidth up to 32, grow stages
inline rsp_status plan_wide(const rsp_chain_params &p, const Regs &r, const Tables &t, WideArgs &a, std::string &err) {
  std::memset(&a, 0, sizeof(a));
  const rsp_fft_params &f = p.fft;
  const int Smax = log2_ceil((uint32_t)f.numPoints);
  const int S = f.runTime ? (int)r.fft[RSP_FFT_REG_FFTSIZE] : Smax;
  if (S < 1 || S > Smax) { err = "fftSize register must be in 1..log2(numPoints)"; return RSP_EINVAL; }
  if ((Smax - S) & 1) { err = "radix 2^2 run-time size must drop an even number of stages (Radix22RunTimeSpec.scala:18-19)"; return RSP_EINVAL; }
  a.S = S; a.w0 = f.dataWidth; a.bp = f.binPoint; a.tw_bp = f.twiddleWidth - 2; a.trim = f.trimType; a.use4muls = f.use4Muls ? 1 : 0;
  a.bitrev = f.useBitReverse ? 1 : 0;
  a.fft_dir = f.fftDirReg ? (int)(r.fft[RSP_FFT_REG_FFTDIR] & 1u) : (f.fftDir ? 1 : 0);
  if (!a.fft_dir && (f.expandLogic != 0 || f.keepMSBorLSBReg)) {
    err = "IFFT with grow stages or keepMSBorLSBReg: the 1/numPoints output scalar of the reference (SDFChainRadix22.scala:362) is not built"; return RSP_EUNSUPPORTED; }
  const uint32_t phys = f.keepMSBorLSBReg ? r.fft[RSP_FFT_REG_KEEPMSBORLSB] : f.keepMSBorLSB;
  const uint32_t smask = (1u << S) - 1u;
  a.keep_mask = (phys >> (Smax - S)) & smask;               // a run-time size uses the LAST S stages (SDFChainRadix22.scala:146-150)
  a.expand_mask = (f.expandLogic >> (Smax - S)) & smask;
  a.base_growth = popcount32(f.expandLogic & ((1u << (Smax - S)) - 1u));
  a.wo = f.dataWidth + a.base_growth + popcount32(a.expand_mask);
  a.out_words = 2 * a.wo > 64 ? 2 : 1;
  for (int L = 0; L < 16; ++L) a.tw_off[L] = t.tw_off[L];
  a.win_on = (f.windowFunc != RSP_WINDOW_NONE && a.fft_dir == 1) ? 1 : 0;
  if (f.windowRam) a.win_on = (r.fft_dowin & 1u) ? 1 : 0;
  a.win_shift = f.windowWidth - 2;
  a.ovf_shift = Smax - S;
  return RSP_OK;
}

// ---- PLFG schedule -----------------------------------------------------------------------------------------------
// The PLFG is a counter tree (frames > chirps > repeats > segments > samples, PLFGDspBlockMem.scala:242-345).  Inside a
// segment the frequency word is an arithmetic progression and the NCO phase a quadratic in the sample index, so the host
// flattens the registers + config RAM into the segments of ONE period (a frame plus the inter-frame pause) with the
// accumulator value in front of each, and the device evaluates any sample in closed form (modulo 2^outputWidth and
// 2^phaseWidth); later periods only shift the accumulator by a constant.
struct GenPlan {
  std::vector<GenSegment> segs;
  uint64_t period_len = 0;     // samples of one frame + pause
  uint64_t frame_len = 0;      // samples of one frame without the pause
  uint32_t period_phase = 0;   // accumulator advance over one period (mod 2^phaseWidth)
  uint64_t total = 0;          // samples the PLFG produces before it stops (~0 = forever)
  // the NCO phase word of every sample of period 0 (phaseWidth <= 16, period up to 2^20 samples; empty otherwise): sample k of period j
  // has phase (phase_tab[k] + j * period_phase) mod 2^phaseWidth, which lets the fused chain kernel generate its own beats at any stride
  std::vector<uint16_t> phase_tab;
};

inline rsp_status plan_generator(const rsp_chain_params &p, const Regs &r, GenPlan &gp, std::string &err) {
  const rsp_plfg_params &g = p.plfg;
  // The config word is sliced at outputWidthInt (point4 of PLFGDspBlockMem.scala:161-171); the slope field and startingPoint (both
  // outputWidthInt bits) are zero-extended into the (Int + Frac)-bit protoOut by asTypeOf and read as a signed raw value.
  const int owi = g.outputWidthInt, ow = g.outputWidthInt + g.outputWidthFrac, nsw = g.maxNumOfSamplesWidth, pw = p.nco.phaseWidth;
  if (pw > ow) { err = "phaseWidth > PLFG output width is not built (closed-form accumulator needs phaseWidth <= outputWidth)"; return RSP_EUNSUPPORTED; }
  const uint32_t pmask = pw >= 32 ? 0xFFFFFFFFu : ((1u << pw) - 1u);
  const long long start = wrap_host((long long)(r.plfg[RSP_PLFG_REG_STARTINGPOINT] & ((1ULL << owi) - 1)), ow);
  const uint32_t frame_num = r.plfg[RSP_PLFG_REG_FRAMENUM];
  const uint32_t inter = r.plfg[RSP_PLFG_REG_INTERFRAMENUMOFSAMPLES];
  const uint32_t ndiff = r.plfg[RSP_PLFG_REG_DIFFERENTCHIRPSNUM] < 1 ? 1 : r.plfg[RSP_PLFG_REG_DIFFERENTCHIRPSNUM];
  const uint32_t seg0 = RSP_PLFG_REG_SEGMENTNUMS0, rep0 = seg0 + g.maxChirpOrdinalNum, ord0 = rep0 + g.maxNumOfDifferentChirps;
  if (ndiff > (uint32_t)g.maxNumOfDifferentChirps) { err = "differentChirpsNum exceeds maxNumOfDifferentChirps"; return RSP_EINVAL; }
  gp.segs.clear(); gp.phase_tab.clear();
  const bool acc_on = p.nco.phaseAccEnable != 0;
  bool tab_ok = pw <= 16;
  auto tab_push = [&](uint32_t acc, long long v) {
    if (!tab_ok) return;
    if (gp.phase_tab.size() >= (1u << 20)) { tab_ok = false; gp.phase_tab.clear(); return; }
    gp.phase_tab.push_back((uint16_t)(acc_on ? acc : ((uint32_t)v & pmask)));
  };
  uint64_t k = 0;
  long long cur = start;
  uint32_t phase = 0;
  for (uint32_t d = 0; d < ndiff; ++d) {
    const uint32_t chirp = r.plfg[ord0 + d];
    const uint32_t reps = r.plfg[rep0 + d] < 1 ? 1 : r.plfg[rep0 + d];   // (cnt + 1) >= n ends the repeat loop (:259)
    if (chirp >= (uint32_t)g.maxChirpOrdinalNum) { err = "chirpOrdinalNums entry out of range"; return RSP_EINVAL; }
    for (uint32_t rep = 0; rep < reps; ++rep) {
      const uint32_t nseg = r.plfg[seg0 + chirp] < 1 ? 1 : r.plfg[seg0 + chirp];
      if (nseg > (uint32_t)g.maxNumOfSegments) { err = "segmentNums entry exceeds maxNumOfSegments"; return RSP_EINVAL; }
      for (uint32_t sg = 0; sg < nseg; ++sg) {
        const uint64_t wd = r.plfg_ram[(size_t)chirp * g.maxNumOfSegments + sg];
        const long long slope = wrap_host((long long)(wd & ((1ULL << owi) - 1)), ow);
        uint32_t ns = (uint32_t)((wd >> owi) & ((1ULL << nsw) - 1));
        if (!(ns > 1)) ns += 1;                                   // PLFGDspBlockMem.scala:175
        const int sign = (int)((wd >> (owi + nsw)) & 1), reset = (int)((wd >> (owi + nsw + 1)) & 1);
        const long long step = sign ? slope : -slope;
        GenSegment s; s.pad_ = 0; s.start = k; s.n = ns; s.phase0 = phase; s.step = (int32_t)step;
        if (reset) { s.f0 = (int32_t)start; s.flags = 1u; }
        else if (sg == 0) { s.f0 = (int32_t)start; s.flags = 0u; }              // first sample of a chirp restarts at startingPoint
        else { s.f0 = (int32_t)wrap_host(cur + step, ow); s.flags = 0u; }
        gp.segs.push_back(s);
        // walk the segment to get the end state exactly as the hardware's wrapping registers do
        for (uint32_t q = 0; q < ns; ++q) {
          const long long v = (s.flags & 1u) ? start : wrap_host((long long)s.f0 + (long long)q * step, ow);
          cur = v;
          phase = (phase + ((uint32_t)v & pmask)) & pmask;
          tab_push(phase, v);
        }
        k += ns;
      }
    }
  }
  gp.frame_len = k;
  // Inter-frame pause, as the RTL counter tree does it (PLFGDspBlockMem.scala:242-275): framePauseCounter starts counting on the tick
  // of the frame's LAST sample and the frame restarts when fpc + 1 >= interframeNumOfSamples, so the stream holds inter - 1 pause
  // samples of startingPoint (none for inter <= 1).  The reference's Python model (plfg_script.py:50-53) emits inter samples; its
  // value-checked tests use frameNum = 1, where both agree.  framePauseCounter is UInt(6.W) (:144): fpc + 1 never reaches 64, so
  // inter >= 64 never leaves the pause in hardware -- refused here when a second frame would be needed.
  const uint32_t pause = inter > 1 ? inter - 1 : 0;
  if (frame_num != 1 && inter > 63) { err = "interframeNumOfSamples > 63 never ends the pause (framePauseCounter is 6 bits wide)"; return RSP_EUNSUPPORTED; }
  if (pause > 0) {
    GenSegment s; s.pad_ = 0; s.start = k; s.n = pause; s.f0 = (int32_t)start; s.step = 0; s.phase0 = phase; s.flags = 1u;
    gp.segs.push_back(s);
    for (uint32_t q = 0; q < pause; ++q) { phase = (phase + ((uint32_t)start & pmask)) & pmask; tab_push(phase, start); }
    k += pause;
  }
  gp.period_len = k;
  gp.period_phase = acc_on ? phase : 0;
  if (!tab_ok || gp.phase_tab.size() != k) gp.phase_tab.clear();
  // frameNum frames with a pause between them, none after the last; 0 = forever
  gp.total = frame_num == 0 ? ~0ULL : (uint64_t)frame_num * gp.period_len - (uint64_t)pause;
  return RSP_OK;
}

}  // namespace rsp
