"""TEST INFRASTRUCTURE (oracle): a pure-Python restatement of the JTAG-to-bus-master path of the reference, bit by bit, to check the
product's rsp_jtag_decode against.  Only tests/ may import this.

 - TAP state machine: generators/jtag2mm/src/main/scala/chisel-jtag/JtagStateMachine.scala:103-149
 - shift registers: chisel-jtag/JtagShifter.scala:136-190 (regs(n-1) := TDI, regs(x) := regs(x+1); capture pattern b01; update = Cat(regs.reverse))
 - controller: jtagToMaster.scala:32-123 (instruction / data latches updated in UpdateIR / UpdateDR, instruction reset in TestLogicReset)
 - bus master: jtagToMaster.scala:530-590 (the `when(currentInstruction === ...)` latches, the edge-triggered should* flags) and the
   address sequences of the burst states (:752, :812)
The one-cycle TMS register of the RTL (JtagStateMachine.scala:95) is transport latency and is left out, as in the product."""

NEXT = {  # state: (next if TMS = 0, next if TMS = 1)
    "TestLogicReset": ("RunTestIdle", "TestLogicReset"), "RunTestIdle": ("RunTestIdle", "SelectDRScan"),
    "SelectDRScan": ("CaptureDR", "SelectIRScan"), "CaptureDR": ("ShiftDR", "Exit1DR"), "ShiftDR": ("ShiftDR", "Exit1DR"),
    "Exit1DR": ("PauseDR", "UpdateDR"), "PauseDR": ("PauseDR", "Exit2DR"), "Exit2DR": ("ShiftDR", "UpdateDR"),
    "UpdateDR": ("RunTestIdle", "SelectDRScan"), "SelectIRScan": ("CaptureIR", "TestLogicReset"),
    "CaptureIR": ("ShiftIR", "Exit1IR"), "ShiftIR": ("ShiftIR", "Exit1IR"), "Exit1IR": ("PauseIR", "UpdateIR"),
    "PauseIR": ("PauseIR", "Exit2IR"), "Exit2IR": ("ShiftIR", "UpdateIR"), "UpdateIR": ("RunTestIdle", "SelectDRScan"),
}


def jtag_decode(samples, ir_length=3, beat_bytes=4, burst_max=8, pins=(0x10, 0x20, 0x40)):
    """-> list of (kind, addr, data), kind 0 = write, 1 = read"""
    tck_m, tms_m, tdi_m = pins
    dr_bits = 8 * beat_bytes
    state = "TestLogicReset"
    ir = [0] * ir_length
    dr = [0] * dr_bits
    instr = prev = 0
    data_out = 0
    data_value = address_value = burst_total = burst_index = 0
    burst_data = [0] * max(1, burst_max)
    ops = []

    def follow():
        nonlocal data_value, address_value, burst_total, burst_index
        if instr == 0b0011: data_value = data_out
        if instr == 0b0010: address_value = (data_out >> (beat_bytes * 4)) & ((1 << (beat_bytes * 4)) - 1)
        if instr == 0b1000: burst_total = (data_out >> ((beat_bytes - 1) * 8)) & 0xFF
        if instr == 0b1010: burst_index = (data_out >> ((beat_bytes - 1) * 8)) & 0xFF
        if instr == 0b1011 and burst_index < burst_max: burst_data[burst_index] = data_out

    def set_instruction(v):
        nonlocal instr, prev
        prev, instr = instr, v
        follow()
        if instr == prev: return
        if instr == 0b0001: ops.append((0, address_value, data_value))
        if instr == 0b0100: ops.append((1, address_value, 0))
        if instr == 0b1001:
            for i in range(min(burst_total, burst_max)): ops.append((0, address_value + i * beat_bytes, burst_data[i]))
        if instr == 0b1100:
            for i in range(burst_total): ops.append((1, address_value + i * beat_bytes, 0))

    last = bool(samples[0] & tck_m) if len(samples) else False
    for s in samples[1:]:
        clk = bool(s & tck_m)
        rising, last = clk and not last, clk
        if not rising: continue
        tms, tdi = int(bool(s & tms_m)), int(bool(s & tdi_m))
        if state == "CaptureIR": ir = [1] + [0] * (ir_length - 1)
        elif state == "ShiftIR": ir = ir[1:] + [tdi]
        elif state == "CaptureDR": dr = [1] + [0] * (dr_bits - 1)
        elif state == "ShiftDR": dr = dr[1:] + [tdi]
        state = NEXT[state][tms]
        if state == "TestLogicReset":
            if instr != 0: set_instruction(0)
        elif state == "UpdateIR":
            set_instruction(sum(b << i for i, b in enumerate(ir)))
        elif state == "UpdateDR":
            data_out = sum(b << i for i, b in enumerate(dr))
            follow()
    return ops


class JtagEncoder:
    """Builds GPIO captures (the inverse of jtag_decode) for tests: the same pin protocol as the reference's gpio_for_jtag.py, plus
    the paths that script never takes (pauses in the middle of a shift, going through Run-Test/Idle, a reset in the middle)."""

    def __init__(self, pins=(0x10, 0x20, 0x40), idle=0):
        self.tck, self.tms, self.tdi = pins
        self.samples = [idle]

    def clock(self, tms, tdi=0, extra=0):
        v = (self.tms if tms else 0) | (self.tdi if tdi else 0) | extra
        self.samples += [v, v | self.tck]

    def reset(self):
        for _ in range(5): self.clock(1)
        self.clock(0)

    def scan(self, value, length, ir, pause_after=None, idle_cycles=0):
        self.clock(1)
        if ir: self.clock(1)
        self.clock(0); self.clock(0)
        for i in range(length):
            last = i == length - 1
            self.clock(1 if last else 0, (value >> i) & 1)
            if pause_after is not None and i == pause_after and not last:
                # that bit went in with TMS = 0; leave through Exit1 -> Pause -> Exit2 -> Shift needs the bit before Exit1 to carry TMS = 1:
                raise ValueError("use scan_with_pause")
        self.clock(1); self.clock(0)
        for _ in range(idle_cycles): self.clock(0)

    def scan_with_pause(self, value, length, ir, split, pause_cycles=2):
        """shift `split` bits, go Exit1 -> Pause (pause_cycles) -> Exit2 -> Shift, shift the rest"""
        assert 0 < split < length
        self.clock(1)
        if ir: self.clock(1)
        self.clock(0); self.clock(0)
        for i in range(split): self.clock(1 if i == split - 1 else 0, (value >> i) & 1)     # last of the first part: -> Exit1
        self.clock(0)                                                                        # Exit1 -> Pause
        for _ in range(pause_cycles - 1): self.clock(0)                                      # stay in Pause
        self.clock(1)                                                                        # Pause -> Exit2
        self.clock(0)                                                                        # Exit2 -> Shift
        for i in range(split, length): self.clock(1 if i == length - 1 else 0, (value >> i) & 1)
        self.clock(1); self.clock(0)

    def bytes(self):
        return bytes(self.samples)
