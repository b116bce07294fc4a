"""Wire formats of the reference's FPGA builds as host-side adapters: a recorded JTAG capture becomes the register
writes it stands for, a UART byte stream becomes stream beats and back.  The decoding itself is the C ABI's (rsp_jtag_decode,
rsp_uart_to_beats, rsp_beats_to_uart in include/rsp_chains.h); this module mirrors the reference's names around it.

 - generators/jtag2mm (JtagController + JTAGToMasterAXI4, jtagToMaster.scala:32-123,525-590; pins as
   generators/jtag2mm/src/test/python/gpio_for_jtag.py:27-52 drives them): `jtag_decode`, `JtagReplay`
 - src/main/scala/RxFftCfarMagTxChain.scala:24-78 (UART RX -> nToOne -> FftMagCfarChainVanilla -> oneToN -> UART TX): `RxFftMagCfarTxChain`"""
from __future__ import annotations

import ctypes as C
from typing import List, NamedTuple

import numpy as np

from . import _abi

TCK, TMS, TDI = 0x10, 0x20, 0x40          # gpio_for_jtag.py:27-52
WRITE, READ = 0, 1


class BusOp(NamedTuple):
    kind: int      # WRITE / READ
    addr: int      # the master's addressValue (16 bits at beatBytes = 4)
    data: int


def jtag_decode(samples: bytes, ir_length: int = 3, beat_bytes: int = 4, burst_max: int = 8,
                pins=(TCK, TMS, TDI)) -> List[BusOp]:
    """GPIO capture -> the AXI4 transactions the JTAG bus master issues (AXI4JTAGToMasterBlock(3, 0, 4, ..., 8) is the reference's
    instantiation, jtagToMaster.scala:914)."""
    samples = bytes(samples)
    n = C.c_uint32(0)
    cap = max(16, len(samples) // 8)
    while True:
        ops = (_abi.BusOp * cap)()
        st = _abi.lib().rsp_jtag_decode(samples, len(samples), pins[0], pins[1], pins[2], ir_length, beat_bytes, burst_max, ops, cap, C.byref(n))
        if st == _abi.RSP_ENOMEM and n.value > cap:
            cap = n.value
            continue
        _abi.check(st)
        return [BusOp(o.kind, o.addr, o.data_lo | (o.data_hi << 32)) for o in ops[:n.value]]


class JtagReplay:
    """Applies a capture to a chain the way the JTAG master would drive its AXI4 bus: memWriteWord(addr_base + addr, data) for a write,
    memReadWord for a read (the values read are returned in order)."""

    def __init__(self, chain, addr_base: int = 0, ir_length: int = 3, burst_max: int = 8):
        self.chain, self.addr_base, self.ir_length, self.burst_max = chain, addr_base, ir_length, burst_max

    def replay(self, samples: bytes) -> List[int]:
        reads = []
        for op in jtag_decode(samples, self.ir_length, self.chain.params.beatBytes, self.burst_max):
            if op.kind == WRITE:
                self.chain.memWriteWord(self.addr_base + op.addr, op.data)
            else:
                reads.append(self.chain.memReadWord(self.addr_base + op.addr))
        return reads


def uart_to_beats(data: bytes, beat_bytes: int = 4, msb_first: bool = False) -> np.ndarray:
    """AXI4StreamWidthAdapter.nToOne(beatBytes): beatBytes received bytes make one beat; a trailing partial beat is not delivered"""
    data = bytes(data)
    out = np.empty(len(data) // beat_bytes, dtype={1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[beat_bytes])
    n = C.c_size_t(0)
    _abi.check(_abi.lib().rsp_uart_to_beats(data, len(data), beat_bytes, int(msb_first), out.ctypes.data, C.byref(n)))
    assert n.value == out.size
    return out


def beats_to_uart(beats: np.ndarray, msb_first: bool = False) -> bytes:
    """AXI4StreamWidthAdapter.oneToN(beatBytes)"""
    beats = np.ascontiguousarray(beats)
    out = C.create_string_buffer(beats.nbytes)
    _abi.check(_abi.lib().rsp_beats_to_uart(beats.ctypes.data, beats.size, beats.itemsize, int(msb_first), C.cast(out, C.c_void_p)))
    return out.raw


class RxFftMagCfarTxChain:
    """src/main/scala/RxFftCfarMagTxChain.scala:24-78: bytes received on the UART are the chain's input beats, its output beats leave
    on the UART.  `paramsChain` as for FftMagCfarChainVanilla; `uartParams` / `divisorInit` describe the serial line itself (baud rate,
    FIFO depths) and change no value, so they are kept for the signature only.  Registers are written through `.chain`
    (memWriteWord, or a JtagReplay around it)."""

    def __init__(self, paramsChain, uartParams=None, divisorInit: int = 0, device: int = 0, msb_first: bool = False):
        from .chain import FftMagCfarChainVanilla
        self.chain = FftMagCfarChainVanilla(paramsChain, device=device)
        self.uartParams, self.divisorInit, self.msb_first = uartParams, divisorInit, msb_first
        self._pending = b""

    def close(self):
        self.chain.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def transfer(self, rx: bytes) -> bytes:
        """feed received bytes, get the bytes transmitted for every chirp they complete (incomplete chirps wait for more input, as the
        FFT's SDF pipeline does)"""
        bb = self.chain.params.beatBytes
        self._pending += bytes(rx)
        chirp_bytes = self.chain.fftSize * bb
        n_chirps = len(self._pending) // chirp_bytes
        if n_chirps == 0:
            return b""
        beats = uart_to_beats(self._pending[:n_chirps * chirp_bytes], bb, self.msb_first)
        self._pending = self._pending[n_chirps * chirp_bytes:]
        out = self.chain.run_host(beats.reshape(n_chirps, self.chain.fftSize), n_chirps, 1)[0]
        return beats_to_uart(np.asarray(out).reshape(-1), self.msb_first)
