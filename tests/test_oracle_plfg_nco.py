"""Pins the oracle's PLFG / NCO restatement (oracle/rsp_oracle.c) against the reference's own known answers:
 - the PLFG RAM words and expected ramps of generators/plfg/src/test/scala/PLFGSpec.scala:42-62
 - outputs of the reference's Python golden PLFG() (tests/golden/plfg_golden.json, made by tests/golden/make_plfg_golden.py)
 - the NCO acceptance criteria of generators/nco/src/test/scala/NCOTester.scala:43-140 (|err| vs sin/cos within tolLSB)
 - calcExpectedNcoOut (src/test/scala/RspChainTesterUtils.scala:174-181)."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def ramp(start, length, slope):           # PLFGTesterUtil.scala:17-28
    return [start + i * slope for i in range(length)]


def triangle(start, l1, l2, s1, s2):      # PLFGTesterUtil.scala:43-59
    out, v = [], start
    for i in range(l1):
        out.append(v)
        if i != l1 - 1:
            v += s1
    for _ in range(l2):
        v -= s2
        out.append(v)
    return out


def test_plfgspec_known_answer(oracle):
    """PLFGSpec.scala:42-62: RAM words 0x1180002, 0x2060000, 0x120000A, 0x0140008 -> ramp ++ constant ++ triangle"""
    p = oracle.Plfg(frame_num=1, interframe_num_of_samples=60, different_chirps_num=1, starting_point=0, segment_nums=[4],
                    repeated_chirp_nums=[1], chirp_ordinal_nums=[0], ram=[0x1180002, 0x2060000, 0x120000A, 0x0140008])
    got = p.run(24 + 6 + 52).tolist()
    want = ramp(0, 24, 2) + [0] * 6 + triangle(10, 32, 20, 10, 8)
    assert got == want


def test_plfg_ram_word_layout(oracle):
    """{reset, sign, nSamples[8], slope[16]} (PLFGDspBlockMem.scala:168-178) reproduces the spec's literals"""
    assert oracle.plfg_ram_word(2, 24, True, False) == 0x1180002
    assert oracle.plfg_ram_word(0, 6, False, True) == 0x2060000
    assert oracle.plfg_ram_word(10, 32, True, False) == 0x120000A
    assert oracle.plfg_ram_word(8, 20, False, False) == 0x0140008


def _golden_cases():
    with open(os.path.join(HERE, "golden", "plfg_golden.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("case", _golden_cases(), ids=lambda c: c["name"])
def test_plfg_matches_reference_python_golden(oracle, case):
    """oracle PLFG == outputs of the reference's plfg_script.PLFG() (exact integers)"""
    n_chirps = case["numOfChirps"]
    ram = [0] * (4 * 4)
    for c in range(n_chirps):
        for s in range(case["numOfSegments"][c]):
            slope = case["slopes"][c][s]
            ram[c * 4 + s] = oracle.plfg_ram_word(abs(slope), case["numOfSamples"][c][s], slope >= 0, case["segmentResets"][c][s])
    p = oracle.Plfg(max_num_of_segments=4, max_chirp_ordinal_num=4, frame_num=case["numOfFrames"],
                    interframe_num_of_samples=case["interFrameNumOfSamples"], different_chirps_num=n_chirps,
                    starting_point=case["startingPoint"] & 0xFFFF, segment_nums=case["numOfSegments"],
                    repeated_chirp_nums=[1] * n_chirps, chirp_ordinal_nums=list(range(n_chirps)), ram=ram, pause_model="script")
    want = case["output"]
    got = p.run(len(want) + 100)
    assert got.tolist() == want        # and the PLFG stops after frameNum frames (no trailing pause)


def test_plfg_bumps_short_segments_and_wraps(oracle):
    """nSamples < 2 is bumped by one (PLFGDspBlockMem.scala:175); the value wraps at outputWidth bits"""
    p = oracle.Plfg(frame_num=1, different_chirps_num=1, starting_point=32000, segment_nums=[2], repeated_chirp_nums=[1],
                    chirp_ordinal_nums=[0], ram=[oracle.plfg_ram_word(500, 1, True, False), oracle.plfg_ram_word(500, 4, True, False)])
    got = p.run(100).tolist()
    assert len(got) == 2 + 4
    assert got[:2] == [32000, 32500]
    assert got[2] == 33000 - 65536     # wrapped to 16 bits


def test_nco_table_entries(oracle):
    """T[i] = round(sin(pi/2 * i / tableSize) * 2^(W-2)) (NCOTable.scala:254-258)"""
    t = oracle.nco_table(128, 16)
    i = np.arange(128)
    want = np.floor(np.sin(i * 0.5 * np.pi / 128) * (1 << 14) + 0.5).astype(np.int64)
    assert (t == want).all()
    assert t[0] == 0 and t.max() < (1 << 14)


@pytest.mark.parametrize("table_size,width,phase_width", [(128, 16, 9), (256, 16, 10), (512, 14, 11), (64, 12, 8)])
def test_nco_against_sin_cos(oracle, table_size, width, phase_width):
    """the reference's NCO acceptance: |out - sin/cos(2 pi phase / 2^pw)| within tolLSB raw LSB for every phase
    (generators/nco/src/test/scala/NCOTester.scala:43-140; phaseWidth = addr bits + 2 -> no interpolation terms)"""
    n = 1 << phase_width
    freq = np.ones(n, np.int32)                        # phase walks 1, 2, ..., 2^pw (wraps to 0)
    words, final = oracle.nco_words(freq, table_size, width, phase_width)
    assert final == 0
    re, im = oracle.unpack_iq(words)
    ph = (np.arange(1, n + 1) % n) * 2.0 * np.pi / n
    scale = 1 << (width - 2)
    assert np.abs(im - np.sin(ph) * scale).max() <= 1.0    # rounding of the table entry only
    assert np.abs(re - np.cos(ph) * scale).max() <= 1.0
    # quadrant symmetries hold exactly
    assert re[n - 1] == scale and im[n - 1] == 0         # phase 0: cos = exactly 1.0 (NCOTable.scala:201,262-263)
    q = n // 4
    assert im[q - 1] == scale and re[q - 1] == 0          # phase pi/2


def test_nco_constant_frequency_gives_expected_bin(oracle):
    """RspChainVanillaTester.scala:85: startingPoint f -> tone at bin f * N / (4 * tableSize); calcExpectedNcoOut
    (RspChainTesterUtils.scala:174-181): sample n carries phase (n + 1) * f"""
    f, n_pts, ts = 16, 1024, 128
    words, _ = oracle.nco_words(np.full(n_pts, f, np.int32), ts, 16, 9)
    re, im = oracle.unpack_iq(words)
    spec = np.abs(np.fft.fft(re + 1j * im))
    assert int(spec.argmax()) == f * n_pts // (4 * ts)
    ph = 2.0 * np.pi * ((np.arange(n_pts) + 1) * f % 512) / 512
    assert np.abs(re - np.cos(ph) * 16384).max() <= 1.0 and np.abs(im - np.sin(ph) * 16384).max() <= 1.0


def test_java_random_known_answer(oracle):
    """java.util.Random(11110L).nextDouble(): the LCG the reference's seeded signals come from.  First values of seed 0
    are the published java.util.Random known answers (0.730967787376657, 0.24053641567148587)."""
    r = oracle.JavaRandom(0)
    assert abs(r.next_double() - 0.730967787376657) < 1e-15
    assert abs(r.next_double() - 0.24053641567148587) < 1e-15


# ---------------------------------------------------------------------------------------------------- RTL counter tree
def _rtl_plfg_ticks(n_out, *, max_seg, nsw, ow, owi, frame_num, inter, ndiff, start_reg, seg_nums, rep_nums, ord_nums, ram):
    """Tick-level restatement of the PLFG counter tree (PLFGDspBlockMem.scala:242-295) with the value path of :298-345 applied in
    emission order: one sample per tick while enable && !end; the pause is the ticks on which every counter sits at its terminal
    value and framePauseCounter (UInt(6.W), truncating +) is above 0.  Written from the Scala, not from the oracle's nested loops."""
    def wrap(v, w):
        v &= (1 << w) - 1
        return v - (1 << w) if v >> (w - 1) else v
    start = wrap(start_reg & ((1 << owi) - 1), ow)
    sc = seg = rep = dc = fc = fpc = 0
    end, cur, out = False, start, []
    while len(out) < n_out and not end:
        chirp = ord_nums[dc]
        wd = ram[chirp * max_seg + seg]
        slope = wrap(wd & ((1 << owi) - 1), ow)
        ns = (wd >> owi) & ((1 << nsw) - 1)
        ns = ns if ns > 1 else ns + 1
        sign, rst = (wd >> (owi + nsw)) & 1, (wd >> (owi + nsw + 1)) & 1
        if fpc > 0 or rst or (seg == 0 and sc == 0):
            cur = start
        else:
            cur = wrap(cur + slope if sign else cur - slope, ow)
        out.append(cur)
        if sc + 1 >= ns:
            if seg + 1 >= seg_nums[chirp]:
                if rep + 1 >= rep_nums[dc]:
                    if dc + 1 >= ndiff:
                        if frame_num != 0 and fc + 1 >= frame_num:
                            end = True
                        elif ((fpc + 1) & 63) >= inter:
                            sc = seg = rep = dc = fpc = 0
                            fc = 0 if frame_num == 0 else fc + 1
                        else:
                            fpc = (fpc + 1) & 63
                    else:
                        sc = seg = rep = 0; dc += 1
                else:
                    sc = seg = 0; rep += 1
            else:
                sc = 0; seg += 1
        else:
            sc += 1
    return out


@pytest.mark.parametrize("frame_num,inter", [(3, 17), (2, 1), (2, 2), (0, 5), (4, 63), (1, 63)])
def test_plfg_pause_follows_the_rtl_counter_tree(oracle, frame_num, inter):
    """multi-frame streams: inter - 1 pause samples between frames (none for inter <= 1), none after the last frame; several chirp
    types through chirpOrdinalNums, repeated chirps, a reset segment"""
    W = oracle.plfg_ram_word
    ram = [0] * 16
    ram[0:2] = [W(3, 20, True, False), W(5, 7, False, False)]                    # chirp 0: up, down
    ram[4:7] = [W(1, 9, True, False), W(0, 4, False, True), W(2, 11, False, False)]   # chirp 1: up, reset, down
    ram[8:9] = [W(7, 1, True, False)]                                            # chirp 2: nSamples 1 is bumped to 2
    cfg = dict(frame_num=frame_num, interframe_num_of_samples=inter, different_chirps_num=3, starting_point=40,
               segment_nums=[2, 3, 1, 1], repeated_chirp_nums=[2, 1, 3], chirp_ordinal_nums=[1, 0, 2], ram=ram)
    n = 700
    got = oracle.Plfg(**cfg).run(n).tolist()
    want = _rtl_plfg_ticks(n, max_seg=4, nsw=8, ow=16, owi=16, frame_num=frame_num, inter=inter, ndiff=3, start_reg=40,
                           seg_nums=[2, 3, 1, 1], rep_nums=[2, 1, 3], ord_nums=[1, 0, 2], ram=ram)
    assert got == want
    frame_len = 2 * (9 + 4 + 11) + (20 + 7) + 3 * 2
    if frame_num == 0:
        assert len(got) == n
    else:
        assert len(got) == min(n, frame_num * frame_len + (frame_num - 1) * max(inter - 1, 0))
    # the script model of the reference's Python golden differs by exactly one sample per pause
    if frame_num not in (0, 1) and inter >= 1:
        script = oracle.Plfg(pause_model="script", **cfg).run(n + 10)
        assert len(script) == min(n + 10, frame_num * frame_len + (frame_num - 1) * inter)


def test_plfg_output_width_frac_slices_at_int_width(oracle):
    """outputWidthFrac > 0 (PLFGDspBlockMem.scala:161-178): the config word is sliced at outputWidthInt, slope and startingPoint are
    zero-extended into the (Int + Frac)-bit protoOut.  Hand-derived: Int 8, Frac 4 -> 12-bit output; slope field 0xF0 = +240 raw (not -16),
    startingPoint 0x90 = +144 raw (not -112): 144, 384, ..., 144 + 240 * 7 = 1824, then 2064 wraps to -2032 at 12 bits"""
    wd = oracle.plfg_ram_word(0xF0, 12, True, False, output_width=8, samples_width=8)
    assert wd == (1 << 16) | (12 << 8) | 0xF0
    p = oracle.Plfg(output_width=12, output_width_int=8, frame_num=1, different_chirps_num=1, starting_point=0x90, segment_nums=[1],
                    repeated_chirp_nums=[1], chirp_ordinal_nums=[0], ram=[wd])
    got = p.run(50).tolist()
    want, v = [], 144
    for k in range(12):
        want.append(v if v < 2048 else v - 4096)
        v = (v + 240) % 4096
    assert got == want
    tick = _rtl_plfg_ticks(50, max_seg=4, nsw=8, ow=12, owi=8, frame_num=1, inter=63, ndiff=1, start_reg=0x90, seg_nums=[1, 1, 1, 1],
                           rep_nums=[1], ord_nums=[0], ram=[wd, 0, 0, 0])
    assert tick == want


def test_nco_output_multiplier(oracle):
    """FixedNCOParams.useMultiplier (NCOLazyModuleBlock.scala:354-364; NCOMultiplierSpec.scala:100 writes the factor 0x1000 = 0.25 in
    Q2.14): sin and cos times the factor, product trimmed back to Q2.14 under roundingMode; with enableMultiplying = 0 the factor is
    1 << 14 = 1.0 and the output is unchanged"""
    freq = np.full(1024, 5, np.int32)
    plain, _ = oracle.nco_words(freq, 256, 16, 10)
    same, _ = oracle.nco_words(freq, 256, 16, 10, use_multiplier=True, enable_multiplying=False, multiplying_factor=0x1000)
    assert (same == plain).all()
    quarter, _ = oracle.nco_words(freq, 256, 16, 10, use_multiplier=True, enable_multiplying=True, multiplying_factor=0x1000)
    pc, ps = oracle.unpack_iq(plain)
    qc, qs = oracle.unpack_iq(quarter)
    assert (qc == (pc + 2) >> 2).all() and (qs == (ps + 2) >> 2).all()          # RoundHalfUp of x / 4
    ph = 2 * np.pi * np.cumsum(freq) / 1024
    assert np.abs(qc - 0.25 * 16384 * np.cos(ph)).max() <= 1.5 and np.abs(qs - 0.25 * 16384 * np.sin(ph)).max() <= 1.5
    neg, _ = oracle.nco_words(freq, 256, 16, 10, use_multiplier=True, enable_multiplying=True, multiplying_factor=0xC000)   # -1.0
    nc, ns = oracle.unpack_iq(neg)
    assert (nc == np.where(pc == 16384, -16384, -pc)).all() and (ns == -ps).all()
    fl, _ = oracle.nco_words(freq, 256, 16, 10, use_multiplier=True, enable_multiplying=True, multiplying_factor=0x1000, rounding_mode="Floor")
    assert (oracle.unpack_iq(fl)[0] == pc >> 2).all()


def test_nco_rasterized_mode():
    """NCOTableRasterizedMode (NCOTable.scala:322-464): address = phase mod tableSize, quadrant = (phase / tableSize) mod 4, phaseWidth =
    log2Ceil(tableSize) + 2 (NCO.scala:56-62).  With a power-of-two table it IS the standard mode; with any other size the 4 * tableSize
    phase values of one turn give sin / cos within half an LSB of the real functions (the NCO specs' tolerance is looser), the reversed look-up of
    address 0 is exactly 1.0, and the phase words past 4 * tableSize start the next turn (the accumulator still wraps at 2^phaseWidth)."""
    from oracle import oracle as O
    f = (np.arange(0, 2000, dtype=np.int32) * 7) % 41
    for ts, pw in ((128, 9), (256, 10), (1024, 12)):
        a, pa = O.nco_words(f, ts, 16, pw, True)
        b, pb = O.nco_words(f, ts, 16, pw, True, rasterized=True)
        assert (a == b).all() and pa == pb
    for ts in (100, 96, 3, 1000, 3000):
        pw = max(1, (ts - 1).bit_length()) + 2
        ph = np.arange(1 << pw, dtype=np.int32)
        w, _ = O.nco_words(ph, ts, 16, pw, False, rasterized=True)
        s = ((w & 0xFFFF).astype(np.int64) ^ 0x8000) - 0x8000
        c = (((w >> 16) & 0xFFFF).astype(np.int64) ^ 0x8000) - 0x8000
        ang = 2 * np.pi * (ph % (4 * ts)) / (4 * ts)
        assert np.abs(s - np.sin(ang) * 16384).max() <= 0.5 + 1e-9 and np.abs(c - np.cos(ang) * 16384).max() <= 0.5 + 1e-9
        assert c[0] == 16384 and s[ts] == 16384 and c[2 * ts] == -16384 and s[3 * ts] == -16384     # the four exact 1.0
        assert (w[4 * ts:] == w[:(1 << pw) - 4 * ts]).all()                                        # next turn before the accumulator wraps
        # accumulating: phase = running sum mod 2^pw, then the same table
        f2 = np.full(500, 37, dtype=np.int32)
        acc, _ = O.nco_words(f2, ts, 16, pw, True, rasterized=True)
        assert (acc == w[(np.cumsum(f2.astype(np.int64)) % (1 << pw))]).all()
