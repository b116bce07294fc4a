"""Wire-format adapters: rsp_jtag_decode / rsp_uart_to_beats / rsp_beats_to_uart of the C ABI against
 - the reference's own JTAG driver script run in the build container (tests/golden/jtag_golden.json, made by make_jtag_golden.py
   from generators/jtag2mm/src/test/python/gpio_for_jtag.py): the register writes its byte sequences stand for
 - oracle/wire.py, the bit-level restatement of JtagStateMachine / JtagShifter / jtagToMaster, on captures that take every path of the
   TAP (pauses, resets, 4-bit instructions with the burst set, reads) and on random pin noise
and, on the GPU, a chain configured through a JTAG capture or fed through the UART framing against the same chain driven directly."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rsp_chains_b200 as R                      # noqa: E402
from rsp_chains_b200 import wire                 # noqa: E402
from oracle import wire as owire                 # noqa: E402


def test_reference_script_sequences_decode_to_its_register_writes():
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "jtag_golden.json")))
    assert (g["pins"]["tck"], g["pins"]["tms"], g["pins"]["tdi"]) == (wire.TCK, wire.TMS, wire.TDI)
    for case in g["cases"]:
        samples = bytes(case["samples"])
        ops = wire.jtag_decode(samples, ir_length=case["ir_length"])
        assert [(o.kind, o.addr, o.data) for o in ops] == [(wire.WRITE, a, d) for a, d in case["writes"]], case["name"]
        assert [tuple(o) for o in ops] == owire.jtag_decode(samples, case["ir_length"])
    # the script's own list: multiply enable, multiplying factor, frequency, NCO enable (gpio_for_jtag.py:58-84)
    assert g["cases"][0]["writes"] == [[0x0000, 1], [0x0004, 0x1000], [0x000C, 8], [0x0008, 1]]


def _burst_capture():
    e = owire.JtagEncoder()
    e.reset()
    want = []
    # single write through a paused data scan, then a read of the same address
    e.scan(0x3, 4, True); e.scan_with_pause(0xCAFEF00D, 32, False, split=13, pause_cycles=3)
    e.scan(0x2, 4, True); e.scan(0x0040, 16, False, idle_cycles=2)
    e.scan(0x1, 4, True); want.append((0, 0x40, 0xCAFEF00D))
    e.scan(0x4, 4, True); want.append((1, 0x40, 0))
    # burst of three writes from 0x100, then a burst read
    e.scan(0x8, 4, True); e.scan(3, 8, False)
    for i, v in enumerate((0x11111111, 0x80000002, 0x7FFFFFF3)):
        e.scan(0xA, 4, True); e.scan(i, 8, False)
        e.scan(0xB, 4, True); e.scan(v, 32, False)
        want.append((0, 0x100 + 4 * i, v))
    e.scan(0x2, 4, True); e.scan(0x0100, 16, False)
    e.scan(0x9, 4, True)
    e.scan(0xC, 4, True); want += [(1, 0x100 + 4 * i, 0) for i in range(3)]
    # the same write instruction twice in a row fires once (edge triggered); after another instruction it fires again
    e.scan(0x1, 4, True); want.append((0, 0x100, 0xCAFEF00D))
    e.scan(0x1, 4, True)
    e.scan(0x3, 4, True); e.scan(5, 32, False); e.scan(0x1, 4, True); want.append((0, 0x100, 5))
    # Test-Logic-Reset puts instruction 0 back: the next write instruction is an edge again
    e.reset(); e.scan(0x1, 4, True); want.append((0, 0x100, 5))
    return e.bytes(), want


def test_every_tap_path_and_the_burst_instructions():
    samples, want = _burst_capture()
    ops = wire.jtag_decode(samples, ir_length=4, burst_max=8)
    assert [tuple(o) for o in ops] == want
    assert owire.jtag_decode(samples, 4) == want
    # other pins, other idle level
    e = owire.JtagEncoder(pins=(0x01, 0x02, 0x80), idle=0x0C)
    e.reset(); e.scan(3, 3, True); e.scan(0xA5A5A5A5, 32, False); e.scan(2, 3, True); e.scan(0x30, 16, False); e.scan(1, 3, True)
    assert [tuple(o) for o in wire.jtag_decode(e.bytes(), 3, pins=(0x01, 0x02, 0x80))] == [(0, 0x30, 0xA5A5A5A5)]
    # beatBytes = 8: 64-bit data register, 32-bit address in its upper half
    e = owire.JtagEncoder()
    e.reset(); e.scan(3, 4, True); e.scan(0x0123456789ABCDEF, 64, False); e.scan(2, 4, True); e.scan(0x30002008, 32, False); e.scan(1, 4, True)
    assert [tuple(o) for o in wire.jtag_decode(e.bytes(), 4, beat_bytes=8)] == [(0, 0x30002008, 0x0123456789ABCDEF)]
    assert owire.jtag_decode(e.bytes(), 4, 8) == [(0, 0x30002008, 0x0123456789ABCDEF)]


def test_random_pin_noise_product_equals_oracle():
    rng = np.random.default_rng(5)
    for trial in range(40):
        ir_length = int(rng.integers(3, 6))
        n = int(rng.integers(200, 6000))
        # TMS biased low so that the walk spends time in the shift states
        tms = rng.random(n) < (0.25 if trial % 2 else 0.45)
        tdi = rng.random(n) < 0.5
        tck = rng.random(n) < 0.5
        samples = bytes((wire.TCK * tck + wire.TMS * tms + wire.TDI * tdi + (rng.integers(0, 16, n))).astype(np.uint8))
        got = [tuple(o) for o in wire.jtag_decode(samples, ir_length, burst_max=4)]
        assert got == owire.jtag_decode(samples, ir_length, 4, 4), trial


def test_decode_argument_errors_and_capacity():
    import ctypes as C
    from rsp_chains_b200 import _abi
    L = _abi.lib()
    n = C.c_uint32(0)
    assert L.rsp_jtag_decode(b"\0", 1, 0x10, 0x20, 0x40, 2, 4, 8, None, 0, C.byref(n)) == _abi.RSP_EINVAL
    assert "irLength" in L.rsp_last_error().decode()
    assert L.rsp_jtag_decode(b"\0", 1, 0x10, 0, 0x40, 3, 4, 8, None, 0, C.byref(n)) == _abi.RSP_EINVAL
    samples, want = _burst_capture()
    ops = (_abi.BusOp * 2)()
    assert L.rsp_jtag_decode(samples, len(samples), 0x10, 0x20, 0x40, 4, 4, 8, ops, 2, C.byref(n)) == _abi.RSP_ENOMEM
    assert n.value == len(want) and (ops[0].addr, ops[0].data_lo) == (0x40, 0xCAFEF00D)
    assert wire.jtag_decode(b"") == []


def test_uart_framing_round_trip():
    rng = np.random.default_rng(2)
    for bb, dt in ((4, np.uint32), (2, np.uint16), (8, np.uint64), (1, np.uint8)):
        beats = rng.integers(0, 1 << min(63, 8 * bb), 257, dtype=np.uint64).astype(dt)
        for msb in (False, True):
            raw = wire.beats_to_uart(beats, msb_first=msb)
            assert len(raw) == beats.size * bb
            first = int(beats[0])
            assert raw[0] == ((first >> (8 * (bb - 1))) & 0xFF if msb else first & 0xFF)
            assert np.array_equal(wire.uart_to_beats(raw, bb, msb), beats)
            # a trailing partial beat is not delivered
            assert np.array_equal(wire.uart_to_beats(raw + b"\x01" * (bb - 1), bb, msb), beats)
    assert np.array_equal(wire.uart_to_beats(bytes([0x78, 0x56, 0x34, 0x12])), np.array([0x12345678], np.uint32))
    assert np.array_equal(wire.uart_to_beats(bytes([0x78, 0x56, 0x34, 0x12]), msb_first=True), np.array([0x78563412], np.uint32))


@pytest.mark.gpu
def test_chain_configured_through_jtag_and_fed_through_uart(cuda):
    """a capture that writes the run-time registers of the vanilla chain through the JTAG master (bus offsets = the low 16 address
    bits of the blocks, FftMagCfarChain.scala app: fft 0x0100, mag 0x0200, cfar 0x2000) gives the same output words as memWriteWord, and
    the chain behind the UART adapters returns the same words as bytes"""
    from rsp_chains_b200 import synth
    n, log_n = 1024, 10
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    regs = [(p.fftAddress.base, log_n), (p.magAddress.base, 2), (p.cfarAddress.base + 0 * 4, n), (p.cfarAddress.base + 1 * 4, 14),
            (p.cfarAddress.base + 2 * 4, 1), (p.cfarAddress.base + 3 * 4, 5), (p.cfarAddress.base + 4 * 4, 0), (p.cfarAddress.base + 6 * 4, 1),
            (p.cfarAddress.base + 7 * 4, 32), (p.cfarAddress.base + 8 * 4, 4)]
    x = synth.chirps_fixed(3, n, seed=4)
    with R.FftMagCfarChainVanilla(p) as ref:
        for a, v in regs: ref.memWriteWord(a, v)
        want = ref.run_host(x, 3, 1)[0]
    base = p.fftAddress.base & ~0xFFFF
    e = owire.JtagEncoder()
    e.reset()
    for a, v in regs:
        e.scan(3, 3, True); e.scan(v, 32, False); e.scan(2, 3, True); e.scan(a - base, 16, False); e.scan(1, 3, True)
    e.scan(2, 3, True); e.scan(p.magAddress.base - base, 16, False); e.scan(4, 3, True)      # read the magnitude selector back
    with R.RxFftMagCfarTxChain(p) as u:
        reads = R.JtagReplay(u.chain, addr_base=base).replay(e.bytes())
        assert reads == [2]
        raw = wire.beats_to_uart(x.reshape(-1))
        tx = u.transfer(raw[:5000]) + u.transfer(raw[5000:])                              # the split falls inside the second chirp
        assert u.chain.kernel_launches >= 1
    assert np.array_equal(wire.uart_to_beats(tx).reshape(3, n), want)
    assert (want & 1).sum() > 0
