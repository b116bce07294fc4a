// rsp_wire.cu -- host-side decoders of the two wire formats the reference's FPGA builds are driven through, so that a
// recorded capture can be replayed into a chain handle.  No device work, no chain state: bytes in, bus operations / stream beats out.
//
//  * JTAG (generators/jtag2mm): a capture is a sequence of GPIO samples, one byte each, with TCK, TMS and TDI on three of its bits
//    (generators/jtag2mm/src/test/python/gpio_for_jtag.py:27-52 writes 0x10 / 0x20 / 0x40).  The samples go through the IEEE 1149.1 TAP
//    state machine (chisel-jtag/JtagStateMachine.scala:103-149), the instruction and data shift registers
//    (chisel-jtag/JtagShifter.scala:136-190: new bit in at the top, capture pattern b01) and the instruction set of the bus master
//    (jtagToMaster.scala:542-590, README "User manual"); what comes out are the AXI4 transactions the master would issue.
//    The RTL registers TMS once more than the standard TAP (JtagStateMachine.scala:95); that cycle of latency is transport timing and is
//    not modelled -- the decoder follows the standard, which is what gpio_for_jtag.py's sequences are written for.
//  * UART (src/main/scala/RxFftCfarMagTxChain.scala:27-46): received bytes are grouped beatBytes at a time into one stream beat
//    (AXI4StreamWidthAdapter.nToOne) and every output beat leaves as beatBytes bytes (oneToN).  The adapter is rocket-dsptools code that is
//    not in the reference tree, so which byte of a beat travels first is a parameter here (least significant first by default).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rsp_chains.h"

extern "C" void rsp_set_last_error_(const char *msg);
static rsp_status wfail(rsp_status st, const std::string &msg) { rsp_set_last_error_(msg.c_str()); return st; }

namespace {

enum Tap : int { TLR, RTI, SEL_DR, CAP_DR, SH_DR, EX1_DR, PAUSE_DR, EX2_DR, UPD_DR, SEL_IR, CAP_IR, SH_IR, EX1_IR, PAUSE_IR, EX2_IR, UPD_IR };

// next state for TMS = 0 / 1 (JtagStateMachine.scala:103-149)
const Tap kNext[16][2] = {
    {RTI, TLR},          // TestLogicReset
    {RTI, SEL_DR},       // RunTestIdle
    {CAP_DR, SEL_IR},    // SelectDRScan
    {SH_DR, EX1_DR},     // CaptureDR
    {SH_DR, EX1_DR},     // ShiftDR
    {PAUSE_DR, UPD_DR},  // Exit1DR
    {PAUSE_DR, EX2_DR},  // PauseDR
    {SH_DR, UPD_DR},     // Exit2DR
    {RTI, SEL_DR},       // UpdateDR
    {CAP_IR, TLR},       // SelectIRScan
    {SH_IR, EX1_IR},     // CaptureIR
    {SH_IR, EX1_IR},     // ShiftIR
    {PAUSE_IR, UPD_IR},  // Exit1IR
    {PAUSE_IR, EX2_IR},  // PauseIR
    {SH_IR, UPD_IR},     // Exit2IR
    {RTI, SEL_DR},       // UpdateIR
};

struct Master {   // the registers of JTAGToMasterAXI4 (jtagToMaster.scala:530-590)
  uint32_t beat_bytes, burst_max;
  uint64_t data_value = 0;
  uint32_t address_value = 0, burst_total = 0, burst_index = 0;
  std::vector<uint64_t> burst_data;
  uint32_t instr = 0, prev_instr = 0;
  uint64_t dr = 0;   // controller.io.dataOut: the data register as last updated
  uint64_t dmask() const { return beat_bytes >= 8 ? ~0ull : ((1ull << (8 * beat_bytes)) - 1ull); }
  // the `when(currentInstruction === ...)` latches, evaluated whenever the instruction or the data register changes
  void follow() {
    switch (instr) {
      case 0x3: data_value = dr; break;
      case 0x2: address_value = (uint32_t)((dr >> (4 * beat_bytes)) & ((1ull << (4 * beat_bytes)) - 1ull)); break;   // UInt((beatBytes * 4).W)
      case 0x8: burst_total = (uint32_t)((dr >> (8 * (beat_bytes - 1))) & 0xffu); break;
      case 0xA: burst_index = (uint32_t)((dr >> (8 * (beat_bytes - 1))) & 0xffu); break;
      case 0xB: if (burst_index < burst_max) burst_data[burst_index] = dr; break;
      default: break;
    }
  }
};

}  // namespace

extern "C" rsp_status rsp_jtag_decode(const uint8_t *samples, size_t n_samples, uint8_t tck_mask, uint8_t tms_mask, uint8_t tdi_mask,
                                      uint32_t ir_length, uint32_t beat_bytes, uint32_t burst_max, rsp_bus_op *ops, uint32_t ops_cap,
                                      uint32_t *n_ops) {
  if (!n_ops || (n_samples && !samples)) return wfail(RSP_EINVAL, "rsp_jtag_decode: null argument");
  if (ir_length < 3 || ir_length > 8) return wfail(RSP_EINVAL, "rsp_jtag_decode: irLength must be 3 .. 8 (require(irLength >= 3), jtagToMaster.scala:33)");
  if (beat_bytes != 4 && beat_bytes != 8) return wfail(RSP_EINVAL, "rsp_jtag_decode: beatBytes must be 4 or 8");
  if (!tck_mask || !tms_mask || !tdi_mask) return wfail(RSP_EINVAL, "rsp_jtag_decode: a pin mask is zero");
  Master m;
  m.beat_bytes = beat_bytes; m.burst_max = burst_max; m.burst_data.assign(burst_max ? burst_max : 1, 0);
  const uint32_t ir_mask = (1u << ir_length) - 1u;
  const int dr_bits = 8 * (int)beat_bytes;
  Tap state = TLR;
  uint32_t ir_shift = 0;
  uint64_t dr_shift = 0;
  uint32_t count = 0;
  bool overflow = false;
  auto emit = [&](uint32_t kind, uint32_t addr, uint64_t data) {
    if (count < ops_cap && ops) { ops[count].kind = kind; ops[count].addr = addr; ops[count].data_lo = (uint32_t)data; ops[count].data_hi = (uint32_t)(data >> 32); }
    else overflow = true;
    ++count;
  };
  auto new_instruction = [&](uint32_t v) {
    m.prev_instr = m.instr; m.instr = v & ir_mask;
    m.follow();
    if (m.instr == m.prev_instr) return;   // `currentInstruction === x && RegNext(currentInstruction) =/= x`: edge triggered
    switch (m.instr) {
      case 0x1: emit(RSP_BUS_WRITE, m.address_value, m.data_value & m.dmask()); break;
      case 0x4: emit(RSP_BUS_READ, m.address_value, 0); break;
      case 0x9: for (uint32_t i = 0; i < m.burst_total && i < m.burst_max; ++i) emit(RSP_BUS_WRITE, m.address_value + i * beat_bytes, m.burst_data[i] & m.dmask()); break;
      case 0xC: for (uint32_t i = 0; i < m.burst_total; ++i) emit(RSP_BUS_READ, m.address_value + i * beat_bytes, 0); break;
      default: break;
    }
  };
  bool tck_prev = n_samples ? (samples[0] & tck_mask) != 0 : false;
  for (size_t i = 1; i < n_samples; ++i) {
    const bool tck = (samples[i] & tck_mask) != 0;
    const bool rising = tck && !tck_prev;
    tck_prev = tck;
    if (!rising) continue;
    const int tms = (samples[i] & tms_mask) ? 1 : 0;
    const uint64_t tdi = (samples[i] & tdi_mask) ? 1 : 0;
    // what the state the edge finds the TAP in does with this edge
    switch (state) {
      case CAP_IR: ir_shift = 0x1u & ir_mask; break;                                                   // capture b01
      case SH_IR: ir_shift = ((ir_shift >> 1) | ((uint32_t)tdi << (ir_length - 1))) & ir_mask; break;  // regs(n-1) := TDI, regs(x) := regs(x+1)
      case UPD_IR: break;
      case CAP_DR: dr_shift = 0x1ull; break;
      case SH_DR: dr_shift = ((dr_shift >> 1) | (tdi << (dr_bits - 1))) & m.dmask(); break;
      default: break;
    }
    state = kNext[state][tms];
    if (state == TLR) { if (m.instr != 0) new_instruction(0); }        // initialInstruction (0 in every instantiation of the reference)
    else if (state == UPD_IR) new_instruction(ir_shift);
    else if (state == UPD_DR) { m.dr = dr_shift; m.follow(); }
  }
  *n_ops = count;
  if (overflow) return wfail(RSP_ENOMEM, "rsp_jtag_decode: more bus operations than ops_cap (n_ops holds the number needed)");
  return RSP_OK;
}

extern "C" rsp_status rsp_uart_to_beats(const uint8_t *bytes, size_t n_bytes, uint32_t beat_bytes, int msb_first, void *beats, size_t *n_beats) {
  if (!n_beats || (n_bytes && (!bytes || !beats))) return wfail(RSP_EINVAL, "rsp_uart_to_beats: null argument");
  if (beat_bytes != 1 && beat_bytes != 2 && beat_bytes != 4 && beat_bytes != 8) return wfail(RSP_EINVAL, "rsp_uart_to_beats: beatBytes must be 1, 2, 4 or 8");
  const size_t nb = n_bytes / beat_bytes;   // a trailing partial beat stays in the adapter, as in the RTL
  uint8_t *o = static_cast<uint8_t *>(beats);
  for (size_t b = 0; b < nb; ++b)
    for (uint32_t k = 0; k < beat_bytes; ++k)   // beats are little-endian words in memory
      o[b * beat_bytes + (msb_first ? beat_bytes - 1 - k : k)] = bytes[b * beat_bytes + k];
  *n_beats = nb;
  return RSP_OK;
}

extern "C" rsp_status rsp_beats_to_uart(const void *beats, size_t n_beats, uint32_t beat_bytes, int msb_first, uint8_t *bytes) {
  if (n_beats && (!bytes || !beats)) return wfail(RSP_EINVAL, "rsp_beats_to_uart: null argument");
  if (beat_bytes != 1 && beat_bytes != 2 && beat_bytes != 4 && beat_bytes != 8) return wfail(RSP_EINVAL, "rsp_beats_to_uart: beatBytes must be 1, 2, 4 or 8");
  const uint8_t *in = static_cast<const uint8_t *>(beats);
  for (size_t b = 0; b < n_beats; ++b)
    for (uint32_t k = 0; k < beat_bytes; ++k)
      bytes[b * beat_bytes + k] = in[b * beat_bytes + (msb_first ? beat_bytes - 1 - k : k)];
  return RSP_OK;
}
