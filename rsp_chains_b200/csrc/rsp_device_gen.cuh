// rsp_device_gen.cuh -- PLFG -> NCO in closed form, one sample per call.
//
// The PLFG is a counter tree whose output inside a segment is an arithmetic progression modulo 2^outputWidth
// (generators/plfg/src/main/scala/PLFGDspBlockMem.scala:298-345); the NCO accumulates it modulo 2^phaseWidth
// (generators/nco/src/main/scala/NCOLazyModuleBlock.scala:313-324) and looks the phase up in a quarter-wave sine table
// (generators/nco/src/main/scala/NCOTable.scala:193-280).  The host flattens registers + config RAM into a list of
// segments with the accumulator value in front of each (rsp_host.hpp, plan_generator); sample k is then
//   value  = f0 + q*step                     (mod 2^outputWidth),   q = k - segment.start
//   phase  = phase0 + (q+1)*f0 + step*q(q+1)/2 (mod 2^phaseWidth)   (phaseWidth <= outputWidth)
// which needs no scan on the device.  The list covers one period (frame + inter-frame pause); period j adds
// j * period_phase to the accumulator.
#pragma once
#include <stdint.h>

#include "rsp_device.cuh"

namespace rsp {

struct GenSegment {
  uint64_t start;     // index of the segment's first sample in the stream
  uint32_t n;         // samples in the segment
  int32_t f0;         // PLFG value of the first sample (raw, sign-extended)
  int32_t step;       // per-sample increment (signed slope; 0 for pauses)
  uint32_t phase0;    // NCO accumulator before the first sample of the segment
  uint32_t flags;     // bit 0: every sample is f0 (reset segment / inter-frame pause)
  uint32_t pad_;
};

struct GenArgs {
  const GenSegment *segs;   // the segments of ONE period (frame + inter-frame pause)
  uint32_t n_segs;
  unsigned long long period_len;   // samples per period
  uint32_t period_phase;           // accumulator advance per period
  const int32_t *nco_tab;   // tableSize entries, T[i] = round(sin(pi/2 * i/tableSize) * 2^(tableWidth-2))
  const uint32_t *nco_full; // or null: the beat Cat(cos16, sin16) of EVERY phase word, 2^phaseWidth entries built by the host with
                            // nco_lookup itself (phaseWidth <= 12): one load instead of two look-ups and the quadrant logic
  int table_size, addr_bits, table_width, phase_width, out_width, phase_acc_enable;
  int rasterized;           // NCOTableRasterizedMode (NCOTable.scala:322-464): address = phase mod tableSize, quadrant = (phase / tableSize) mod 4
  uint32_t poff;
  // output multiplier (NCOLazyModuleBlock.scala:354-364): mult_on = the factor is not exactly 1.0, mult_f = factor as FixedPoint(w, w - 2) raw
  int mult_on, mult_f, mult_trim;
};

// sin / cos times the multiplier's factor: context_* with binaryPointGrowth 0 under roundingMode, reinterpreted as protoOut
RSP_HD uint32_t nco_multiply(const GenArgs &g, uint32_t beat) {
  const int w = g.table_width;
  const int co = wrapw((int)(beat >> 16), w), s = wrapw((int)(beat & 0xFFFFu), w);
  const int mc = wrapw((int)trim_rt<long long>((long long)co * g.mult_f, w - 2, g.mult_trim), w);
  const int ms = wrapw((int)trim_rt<long long>((long long)s * g.mult_f, w - 2, g.mult_trim), w);
  return (((uint32_t)mc & 0xFFFFu) << 16) | ((uint32_t)ms & 0xFFFFu);
}

// Cat(cos16, sin16) of one phase word (NCOTable.scala:193-235; NCOLazyModuleBlock.scala:452)
RSP_HD uint32_t nco_lookup(const GenArgs &g, uint32_t ph) {
  const int pw = g.phase_width, ab = g.addr_bits, ts = g.table_size;
  uint32_t msbs, addr, raddr;
  if (g.rasterized) {   // tableSize need not be a power of two; the reversed address tableSize - 0 is outside the table and stands for 1.0 (:362-364)
    addr = ph % (uint32_t)ts; msbs = (ph / (uint32_t)ts) & 3u; raddr = addr == 0 ? 0u : (uint32_t)ts - addr;
  } else {
    msbs = ph >> (pw - 2);
    addr = (ph >> (pw - 2 - ab)) & (uint32_t)(ts - 1);
    raddr = (0u - addr) & (uint32_t)(ts - 1);
  }
  const int one = 1 << (g.table_width - 2);
  const int fwd = ldg(&g.nco_tab[addr]);
  const int rev = addr == 0 ? one : ldg(&g.nco_tab[raddr]);   // the reversed look-up of address 0 is exactly 1.0 (:201,211,262-263)
  int s, co;
  switch (msbs) {
  case 0: s = fwd; co = rev; break;
  case 1: s = rev; co = -fwd; break;
  case 2: s = -fwd; co = -rev; break;
  default: s = -rev; co = fwd; break;
  }
  return (((uint32_t)co & 0xFFFFu) << 16) | ((uint32_t)s & 0xFFFFu);
}

// sample k of the stream: returns the NCO beat, *plfg_value = the PLFG word (sign-extended)
RSP_HD uint32_t gen_sample(const GenArgs &g, unsigned long long k, int *plfg_value) {
  const unsigned long long per = k / g.period_len;
  k -= per * g.period_len;
  // last segment with start <= k
  uint32_t lo = 0, hi = g.n_segs;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (ldg(&g.segs[mid].start) <= k) lo = mid; else hi = mid;
  }
  const GenSegment sg = g.segs[lo];
  const unsigned long long q = k - sg.start;
  const bool constant = (sg.flags & 1u) != 0;
  const long long vq = constant ? (long long)sg.f0 : (long long)sg.f0 + (long long)q * (long long)sg.step;
  const int v = (int)wrapw64(vq, g.out_width);
  *plfg_value = v;
  const uint32_t pmask = g.phase_width >= 32 ? 0xFFFFFFFFu : ((1u << g.phase_width) - 1u);
  uint32_t phase;
  if (g.phase_acc_enable) {
    unsigned long long acc = (unsigned long long)sg.phase0 + per * (unsigned long long)g.period_phase + (q + 1ULL) * (unsigned long long)(long long)sg.f0;
    if (!constant) acc += (unsigned long long)(long long)sg.step * ((q * (q + 1ULL)) >> 1);
    phase = (uint32_t)acc & pmask;
  } else {
    phase = (uint32_t)v & pmask;
  }
  const uint32_t beat = nco_lookup(g, (phase + g.poff) & pmask);
  return g.mult_on ? nco_multiply(g, beat) : beat;
}

// Running form for a thread that produces consecutive samples: the first one in closed form (gen_start), the following ones by
// the recurrences value += step, accumulator += value while the segment lasts (gen_next), all modulo 2^32, which contains the
// moduli 2^outputWidth and 2^phaseWidth (phaseWidth <= outputWidth <= 16); a sample that opens a new segment (or period) is
// evaluated in closed form again.  Same bits as gen_sample, without its 64-bit division and segment search per sample.
struct GenState {
  uint32_t left;      // samples of the current segment after this one
  uint32_t v;         // PLFG value of this sample, unwrapped (mod 2^32)
  uint32_t acc;       // NCO accumulator including this sample (mod 2^32)
  uint32_t step;      // 0 in constant segments
};

RSP_HD void gen_start(const GenArgs &g, unsigned long long k, GenState &st) {
  unsigned long long per;
  if (((k | g.period_len) >> 32) == 0) per = (uint32_t)k / (uint32_t)g.period_len;   // the common case without the 64-bit division
  else per = k / g.period_len;
  k -= per * g.period_len;
  uint32_t lo = 0, hi = g.n_segs;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (ldg(&g.segs[mid].start) <= k) lo = mid; else hi = mid;
  }
  const GenSegment sg = g.segs[lo];
  const uint32_t q = (uint32_t)(k - sg.start);
  const bool constant = (sg.flags & 1u) != 0;
  st.step = constant ? 0u : (uint32_t)sg.step;
  st.left = sg.n - 1u - q;
  st.v = (uint32_t)sg.f0 + q * st.step;
  const uint32_t tri = (uint32_t)(((unsigned long long)q * ((unsigned long long)q + 1ULL)) >> 1);
  st.acc = sg.phase0 + (uint32_t)per * g.period_phase + (q + 1u) * (uint32_t)sg.f0 + st.step * tri;
}
RSP_HD void gen_next(const GenArgs &g, unsigned long long k, GenState &st) {   // k = index of the sample stepped TO
  if (st.left != 0u) { --st.left; st.v += st.step; st.acc += st.v; }
  else gen_start(g, k, st);
}
RSP_HD uint32_t gen_beat(const GenArgs &g, const GenState &st, int *plfg_value) {
  const int v = wrapw((int)st.v, g.out_width);
  *plfg_value = v;
  const uint32_t pmask = g.phase_width >= 32 ? 0xFFFFFFFFu : ((1u << g.phase_width) - 1u);
  const uint32_t phase = (g.phase_acc_enable ? st.acc : (uint32_t)v) & pmask;
  const uint32_t ph = (phase + g.poff) & pmask;
  const uint32_t beat = g.nco_full != nullptr ? ldg(&g.nco_full[ph]) : nco_lookup(g, ph);
  return g.mult_on ? nco_multiply(g, beat) : beat;
}

}  // namespace rsp
