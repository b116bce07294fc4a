"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product package (rsp_chains_b200) never does.  See oracle/rsp_oracle.h for the parity status
("parity unpinned" at the bit level; tolerance- and KAT-pinned).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "librsp_oracle.so")

TRIM = {
    "RoundHalfUp": 0, "Floor": 1, "RoundDown": 1, "Convergent": 2, "RoundHalfToEven": 2, "RoundUp": 3,
    "RoundTowardsZero": 4, "RoundTowardsInfinity": 5, "RoundHalfDown": 6, "RoundHalfTowardsZero": 7,
    "RoundHalfTowardsInfinity": 8, "RoundHalfToOdd": 9, "NoTrim": 10,
}
CORE_MEM, CORE_LIS, CORE_ASR, CORE_OS = 0, 1, 2, 3


def build(force: bool = False) -> str:
    """Compile librsp_oracle.so with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "rsp_oracle.c")
    hdr = os.path.join(_HERE, "rsp_oracle.h")
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_LIB_PATH) for p in (src, hdr))
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _LIB_PATH


class FftCfg(C.Structure):
    _fields_ = [("log_n", C.c_int32), ("data_width", C.c_int32), ("bin_point", C.c_int32),
                ("twiddle_width", C.c_int32), ("trim_type", C.c_int32), ("use4muls", C.c_int32),
                ("bit_reverse", C.c_int32), ("keep_msb_mask", C.c_uint32), ("fft_dir", C.c_int32),
                ("sdf_radix", C.c_int32), ("decim_dit", C.c_int32), ("num_stages", C.c_int32),
                ("expand_mask", C.c_uint32), ("base_growth", C.c_int32)]


class MagCfg(C.Structure):
    _fields_ = [("data_width", C.c_int32), ("bin_point", C.c_int32), ("data_width_log", C.c_int32),
                ("bin_point_log", C.c_int32), ("log2_lookup_width", C.c_int32), ("sel", C.c_int32),
                ("ctx_trim", C.c_int32)]


class CfarCfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "core", "in_width", "in_bp", "thr_width", "thr_bp", "scaler_width", "scaler_bp",
        "lead_lagg_window_size", "guard_window_size", "fft_size", "min_sub_window_size", "send_cut",
        "fft_win", "threshold_scaler", "log_or_linear", "div_sum", "peak_grouping", "cfar_algorithm",
        "cfar_mode", "window_cells", "guard_cells", "index_lagg", "index_lead", "sub_cells", "ctx_trim",
        "edge_mode")]


class ChainCfg(C.Structure):
    _fields_ = [("fft", FftCfg), ("mag", MagCfg), ("cfar", CfarCfg)]


class PlfgCfg(C.Structure):
    _fields_ = [("max_num_of_segments", C.c_int32), ("max_num_of_samples_width", C.c_int32),
                ("output_width", C.c_int32), ("frame_num", C.c_int32), ("interframe_num_of_samples", C.c_int32),
                ("different_chirps_num", C.c_int32), ("starting_point", C.c_int32),
                ("segment_nums", C.POINTER(C.c_uint32)), ("repeated_chirp_nums", C.POINTER(C.c_uint32)),
                ("chirp_ordinal_nums", C.POINTER(C.c_uint32)), ("ram", C.POINTER(C.c_uint32)),
                ("output_width_int", C.c_int32), ("pause_model", C.c_int32)]


class NcoCfg(C.Structure):
    _fields_ = [("table_size", C.c_int32), ("table_width", C.c_int32), ("phase_width", C.c_int32),
                ("phase_acc_enable", C.c_int32), ("phase0", C.c_uint32), ("poff", C.c_uint32),
                ("use_multiplier", C.c_int32), ("enable_multiplying", C.c_int32), ("multiplying_factor", C.c_uint32),
                ("rounding_mode", C.c_int32), ("rasterized", C.c_int32)]


class JRand(C.Structure):
    _fields_ = [("seed", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.rspo_trim.restype = C.c_int64
        L.rspo_trim.argtypes = [C.c_int64, C.c_int, C.c_int]
        L.rspo_wrap.restype = C.c_int64
        L.rspo_wrap.argtypes = [C.c_int64, C.c_int]
        L.rspo_jpl_f64.restype = C.c_double
        L.rspo_jpl_f64.argtypes = [C.c_double, C.c_double]
        L.rspo_jrand_next_double.restype = C.c_double
        L.rspo_plfg.restype = C.c_int64
        L.rspo_nco_words.restype = C.c_uint32
        L.rspo_mag_one.restype = C.c_int32
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


# ---------------------------------------------------------------------------------------------- configs

def fft_cfg(log_n, data_width=16, bin_point=14, twiddle_width=16, trim="RoundHalfUp", use4muls=False,
            bit_reverse=True, keep_msb_mask=0xFFFFFFFF, fft_dir=True, sdf_radix="2^2", decim_dit=False, num_stages=0,
            expand_mask=0, base_growth=0) -> FftCfg:
    """sdf_radix "2" selects SDFChainRadix2 (DIF or DIT); "2^2" SDFChainRadix22 (DIF only)"""
    return FftCfg(log_n, data_width, bin_point, twiddle_width, TRIM[trim] if isinstance(trim, str) else trim,
                  int(use4muls), int(bit_reverse), keep_msb_mask & 0xFFFFFFFF, int(fft_dir),
                  2 if str(sdf_radix) == "2" else 4, int(decim_dit), int(num_stages), expand_mask & 0xFFFFFFFF, int(base_growth))


def mag_cfg(data_width=16, bin_point=14, data_width_log=16, bin_point_log=9, log2_lookup_width=9, sel=2,
            ctx_trim="Floor") -> MagCfg:
    return MagCfg(data_width, bin_point, data_width_log, bin_point_log, log2_lookup_width, sel,
                  TRIM[ctx_trim] if isinstance(ctx_trim, str) else ctx_trim)


def cfar_cfg(fft_win, window_cells, guard_cells, threshold_scaler, *, core=CORE_MEM, in_width=16, in_bp=0,
             thr_width=16, thr_bp=None, scaler_width=16, scaler_bp=0, lead_lagg_window_size=None,
             guard_window_size=None, fft_size=None, min_sub_window_size=0, send_cut=False, log_or_linear=1,
             div_sum=None, peak_grouping=0, cfar_algorithm=0, cfar_mode=0, index_lagg=1, index_lead=1,
             sub_cells=None, ctx_trim="Floor") -> CfarCfg:
    if thr_bp is None:
        thr_bp = in_bp
    if lead_lagg_window_size is None:
        lead_lagg_window_size = window_cells
    if guard_window_size is None:
        guard_window_size = guard_cells
    if fft_size is None:
        fft_size = fft_win
    if div_sum is None:
        div_sum = int(np.log2(window_cells))
    if sub_cells is None:
        sub_cells = window_cells
    return CfarCfg(core, in_width, in_bp, thr_width, thr_bp, scaler_width, scaler_bp, lead_lagg_window_size,
                   guard_window_size, fft_size, min_sub_window_size, int(send_cut), fft_win, threshold_scaler,
                   log_or_linear, div_sum, peak_grouping, cfar_algorithm, cfar_mode, window_cells, guard_cells,
                   index_lagg, index_lead, sub_cells, TRIM[ctx_trim] if isinstance(ctx_trim, str) else ctx_trim, 0)


# ---------------------------------------------------------------------------------------------- tables

def twiddle_table(stage_n, tw_width=16):
    n = 3 * stage_n // 4
    re = np.zeros(max(n, 1), np.int32)
    im = np.zeros(max(n, 1), np.int32)
    lib().rspo_twiddle_table(stage_n, tw_width, _p(re, C.c_int32), _p(im, C.c_int32))
    return re[:n], im[:n]


def twiddle_table_radix2(delay, tw_width=16):
    """stage ROM of the radix-2 chain (SDFChainRadix2.scala:300-306)"""
    re = np.zeros(delay, np.int32)
    im = np.zeros(delay, np.int32)
    lib().rspo_twiddle_table_radix2(delay, tw_width, _p(re, C.c_int32), _p(im, C.c_int32))
    return re, im


def nco_table(table_size, table_width):
    t = np.zeros(table_size, np.int32)
    lib().rspo_nco_table(table_size, table_width, _p(t, C.c_int32))
    return t


def log2_table(lookup_width, bp_log):
    t = np.zeros(1 << lookup_width, np.int32)
    lib().rspo_log2_table(lookup_width, bp_log, _p(t, C.c_int32))
    return t


# ---------------------------------------------------------------------------------------------- FFT

def fft_fixed(cfg: FftCfg, re, im):
    re, im = _i32(re), _i32(im)
    n = 1 << cfg.log_n
    assert re.shape == (n,) and im.shape == (n,)
    ore, oim = np.zeros(n, np.int32), np.zeros(n, np.int32)
    rc = lib().rspo_fft_fixed(C.byref(cfg), _p(re, C.c_int32), _p(im, C.c_int32), _p(ore, C.c_int32), _p(oim, C.c_int32))
    if rc:
        raise ValueError(f"rspo_fft_fixed rc={rc}")
    return ore, oim


def fft_fixed_ex(cfg: FftCfg, re, im):
    """as fft_fixed, plus the per-stage overflow mask (SDFChainRadix22.scala:473-486)"""
    re, im = _i32(re), _i32(im)
    n = 1 << cfg.log_n
    ore, oim = np.zeros(n, np.int32), np.zeros(n, np.int32)
    m = C.c_uint32(0)
    rc = lib().rspo_fft_fixed_ex(C.byref(cfg), _p(re, C.c_int32), _p(im, C.c_int32), _p(ore, C.c_int32), _p(oim, C.c_int32), C.byref(m))
    if rc:
        raise ValueError(f"rspo_fft_fixed_ex rc={rc}")
    return ore, oim, m.value


def fft_stage_width(cfg: FftCfg, stage):
    return int(lib().rspo_fft_stage_width(C.byref(cfg), int(stage)))


def fft_fixed_words64(cfg: FftCfg, words):
    """64-bit beats Cat(re, im) (wide data / grow stages); returns (beats, overflow mask)"""
    words = np.ascontiguousarray(words, dtype=np.uint64)
    n = 1 << cfg.log_n
    flat = words.reshape(-1)
    assert flat.size % n == 0
    wo = fft_stage_width(cfg, cfg.log_n - 1)
    k = 2 if 2 * wo > 64 else 1          # 128-bit beats: two words per beat, low word first
    out = np.zeros(flat.size * k, np.uint64)
    m = C.c_uint32(0)
    rc = lib().rspo_fft_fixed_words64(C.byref(cfg), _p(flat, C.c_uint64), _p(out, C.c_uint64), flat.size // n, C.byref(m))
    if rc:
        raise ValueError(f"rspo_fft_fixed_words64 rc={rc}")
    return (out.reshape(words.shape) if k == 1 else out.reshape(words.shape + (2,))), m.value


def fft_fixed_words_ovf(cfg: FftCfg, words):
    words = _u32(words)
    n = 1 << cfg.log_n
    flat = words.reshape(-1)
    out = np.zeros_like(flat)
    m = C.c_uint32(0)
    rc = lib().rspo_fft_fixed_words_ovf(C.byref(cfg), _p(flat, C.c_uint32), _p(out, C.c_uint32), flat.size // n, C.byref(m))
    if rc:
        raise ValueError(f"rspo_fft_fixed_words_ovf rc={rc}")
    return out.reshape(words.shape), m.value


def fft_f64(log_n, re, im, bit_reverse=True):
    re, im = _f64(re), _f64(im)
    n = 1 << log_n
    ore, oim = np.zeros(n), np.zeros(n)
    rc = lib().rspo_fft_f64(log_n, int(bit_reverse), _p(re, C.c_double), _p(im, C.c_double), _p(ore, C.c_double),
                            _p(oim, C.c_double))
    if rc:
        raise ValueError(f"rspo_fft_f64 rc={rc}")
    return ore, oim


def fft_fixed_words(cfg: FftCfg, words):
    words = _u32(words)
    n = 1 << cfg.log_n
    flat = words.reshape(-1)
    assert flat.size % n == 0
    out = np.zeros_like(flat)
    rc = lib().rspo_fft_fixed_words(C.byref(cfg), _p(flat, C.c_uint32), _p(out, C.c_uint32), flat.size // n)
    if rc:
        raise ValueError(f"rspo_fft_fixed_words rc={rc}")
    return out.reshape(words.shape)


# ---------------------------------------------------------------------------------------------- mag

def mag_words(cfg: MagCfg, words):
    words = _u32(words)
    flat = words.reshape(-1)
    out = np.zeros_like(flat)
    rc = lib().rspo_mag_words(C.byref(cfg), _p(flat, C.c_uint32), _p(out, C.c_uint32), C.c_int64(flat.size))
    if rc:
        raise ValueError(f"rspo_mag_words rc={rc}")
    return out.reshape(words.shape)


def jpl_f64(re, im):
    return lib().rspo_jpl_f64(float(re), float(im))


# ---------------------------------------------------------------------------------------------- CFAR

def cfar_fixed(cfg: CfarCfg, x):
    x = _i32(x)
    n = cfg.fft_win
    assert x.shape == (n,)
    thr_full = np.zeros(n, np.int64)
    thr = np.zeros(n, np.int32)
    peak = np.zeros(n, np.uint8)
    rc = lib().rspo_cfar_fixed(C.byref(cfg), _p(x, C.c_int32), _p(thr_full, C.c_int64), _p(thr, C.c_int32),
                               _p(peak, C.c_uint8))
    if rc:
        raise ValueError(f"rspo_cfar_fixed rc={rc}")
    return thr_full, thr, peak


def cfar_words(cfg: CfarCfg, words):
    words = _u32(words)
    flat = words.reshape(-1)
    assert flat.size % cfg.fft_win == 0
    out = np.zeros_like(flat)
    rc = lib().rspo_cfar_words(C.byref(cfg), _p(flat, C.c_uint32), _p(out, C.c_uint32), flat.size // cfg.fft_win)
    if rc:
        raise ValueError(f"rspo_cfar_words rc={rc}")
    return out.reshape(words.shape)


def cfar_f64(cfg: CfarCfg, scaler, x):
    x = _f64(x)
    thr = np.zeros(cfg.fft_win)
    peak = np.zeros(cfg.fft_win, np.uint8)
    rc = lib().rspo_cfar_f64(C.byref(cfg), C.c_double(scaler), _p(x, C.c_double), _p(thr, C.c_double), _p(peak, C.c_uint8))
    if rc:
        raise ValueError(f"rspo_cfar_f64 rc={rc}")
    return thr, peak


def _golden(fn, sig, *args):
    sig = _f64(sig)
    n = sig.size
    thr = np.zeros(3 * n + 16)
    peaks = np.zeros(2 * n + 16, np.int32)
    nt, npk = C.c_int(0), C.c_int(0)
    rc = fn(_p(sig, C.c_double), n, *args, _p(thr, C.c_double), C.byref(nt), _p(peaks, C.c_int32), C.byref(npk))
    if rc:
        raise ValueError(f"golden rc={rc}")
    return thr[:nt.value].copy(), peaks[:npk.value].copy()


_MODES = {"Cell Averaging": 0, "Greatest Of": 1, "Smallest Of": 2}


def golden_cfar_ca(sig, cfar_mode, reference_cells, guard_cells, consider_edges=False, scaling_factor=1.0,
                   log_mode=False):
    """CFARUtils.cfarCA (generators/cfar/src/main/scala/CFARUtils.scala:84-156)."""
    return _golden(lib().rspo_golden_cfar_ca, sig, _MODES[cfar_mode], reference_cells, guard_cells,
                   int(consider_edges), C.c_double(scaling_factor), int(log_mode))


def golden_cfar_os(sig, reference_cells, guard_cells, scaling_factor=1.0, log_mode=False):
    """CFARUtils.cfarOS (CFARUtils.scala:159-196)."""
    return _golden(lib().rspo_golden_cfar_os, sig, reference_cells, guard_cells, C.c_double(scaling_factor),
                   int(log_mode))


def golden_cfar_gos(sig, reference_cells, guard_cells, index_lagg, index_lead, cfar_mode, consider_edges=False,
                    scaling_factor=1.0, log_mode=False):
    """CFARUtils.cfarGOS (CFARUtils.scala:199-272)."""
    return _golden(lib().rspo_golden_cfar_gos, sig, reference_cells, guard_cells, index_lagg, index_lead,
                   _MODES[cfar_mode], int(consider_edges), C.c_double(scaling_factor), int(log_mode))


def golden_cfar_cash(sig, reference_cells, sub_cells, guard_cells=0, scaling_factor=1.0, log_mode=False):
    """CFARUtils.cfarCASH (CFARUtils.scala:29-81)."""
    return _golden(lib().rspo_golden_cfar_cash, sig, reference_cells, sub_cells, guard_cells,
                   C.c_double(scaling_factor), int(log_mode))


# ---------------------------------------------------------------------------------------------- PLFG / NCO

@dataclass
class Plfg:
    """Run-time state of PLFGDspBlockMem (registers + config RAM), PLFGDspBlockMem.scala:182-217."""
    max_num_of_segments: int = 4
    max_num_of_samples_width: int = 8
    output_width: int = 16
    max_chirp_ordinal_num: int = 4
    max_num_of_different_chirps: int = 8
    frame_num: int = 1
    interframe_num_of_samples: int = 63
    different_chirps_num: int = 1
    starting_point: int = 0
    segment_nums: list = field(default_factory=list)
    repeated_chirp_nums: list = field(default_factory=list)
    chirp_ordinal_nums: list = field(default_factory=list)
    ram: list = field(default_factory=list)
    output_width_int: int = 0       # outputWidthInt when outputWidthFrac > 0 (0 = output_width): the RAM word is sliced there
    pause_model: str = "rtl"        # "rtl": inter - 1 pause samples (PLFGDspBlockMem.scala:242-275); "script": inter (plfg_script.py:50-53)

    def run(self, n_out):
        seg = np.ones(self.max_chirp_ordinal_num, np.uint32)
        seg[:len(self.segment_nums)] = self.segment_nums
        rep = np.zeros(self.max_num_of_different_chirps, np.uint32)
        rep[:len(self.repeated_chirp_nums)] = self.repeated_chirp_nums
        ordn = np.zeros(self.max_num_of_different_chirps, np.uint32)
        ordn[:len(self.chirp_ordinal_nums)] = self.chirp_ordinal_nums
        ram = np.zeros(self.max_chirp_ordinal_num * self.max_num_of_segments, np.uint32)
        ram[:len(self.ram)] = self.ram
        cfg = PlfgCfg(self.max_num_of_segments, self.max_num_of_samples_width, self.output_width, self.frame_num,
                      self.interframe_num_of_samples, self.different_chirps_num, self.starting_point,
                      _p(seg, C.c_uint32), _p(rep, C.c_uint32), _p(ordn, C.c_uint32), _p(ram, C.c_uint32),
                      self.output_width_int, 1 if self.pause_model == "script" else 0)
        out = np.zeros(n_out, np.int32)
        k = lib().rspo_plfg(C.byref(cfg), _p(out, C.c_int32), C.c_int64(n_out))
        return out[:k]


def plfg_ram_word(slope, n_samples, sign_plus, reset, output_width=16, samples_width=8):
    """{reset, sign, nSamples, slope} (PLFGDspBlockMem.scala:168-178)."""
    return ((int(reset) << (output_width + samples_width + 1)) | (int(sign_plus) << (output_width + samples_width))
            | ((n_samples & ((1 << samples_width) - 1)) << output_width) | (slope & ((1 << output_width) - 1)))


def nco_words(freq, table_size=128, table_width=16, phase_width=9, phase_acc_enable=True, phase0=0, poff=0, use_multiplier=False,
              enable_multiplying=False, multiplying_factor=0, rounding_mode="RoundHalfUp", rasterized=False):
    freq = _i32(freq)
    out = np.zeros(freq.size, np.uint32)
    cfg = NcoCfg(table_size, table_width, phase_width, int(phase_acc_enable), phase0, poff, int(use_multiplier), int(enable_multiplying),
                 int(multiplying_factor) & 0xFFFFFFFF, TRIM[rounding_mode] if isinstance(rounding_mode, str) else rounding_mode, int(rasterized))
    ph = lib().rspo_nco_words(C.byref(cfg), _p(freq, C.c_int32), _p(out, C.c_uint32), C.c_int64(freq.size))
    return out, ph


# ---------------------------------------------------------------------------------------------- chain

def chain_fixed_words(fft: FftCfg, mag: MagCfg, cfar: CfarCfg, words, n_threads=1, want_stages=False):
    words = _u32(words)
    n = 1 << fft.log_n
    flat = words.reshape(-1)
    assert flat.size % n == 0
    out = np.zeros_like(flat)
    f = np.zeros_like(flat) if want_stages else None
    m = np.zeros_like(flat) if want_stages else None
    cfg = ChainCfg(fft, mag, cfar)
    rc = lib().rspo_chain_fixed_words(C.byref(cfg), _p(flat, C.c_uint32), _p(out, C.c_uint32), flat.size // n,
                                      int(n_threads), _p(f, C.c_uint32) if want_stages else None,
                                      _p(m, C.c_uint32) if want_stages else None)
    if rc:
        raise ValueError(f"rspo_chain_fixed_words rc={rc}")
    if want_stages:
        return out.reshape(words.shape), f.reshape(words.shape), m.reshape(words.shape)
    return out.reshape(words.shape)


def max_threads():
    return int(lib().rspo_max_threads())


# ---------------------------------------------------------------------------------------------- helpers

class JavaRandom:
    """java.util.Random / scala.util.Random (used by the reference's seeded test signals)."""

    def __init__(self, seed):
        self._r = JRand()
        lib().rspo_jrand_seed(C.byref(self._r), C.c_int64(seed))

    def next_double(self):
        return lib().rspo_jrand_next_double(C.byref(self._r))


def pack_iq(re, im):
    re = np.asarray(re).astype(np.int64)
    im = np.asarray(im).astype(np.int64)
    return (((re & 0xFFFF) << 16) | (im & 0xFFFF)).astype(np.uint32)


def unpack_iq(words):
    w = np.asarray(words, dtype=np.uint32)
    re = (w >> 16).astype(np.uint16).astype(np.int16).astype(np.int32)
    im = (w & 0xFFFF).astype(np.uint16).astype(np.int16).astype(np.int32)
    return re, im


def trim(v, shift, mode):
    return int(lib().rspo_trim(C.c_int64(int(v)), shift, TRIM[mode] if isinstance(mode, str) else mode))


# ---------------------------------------------------------------------------------------------------------------------
# window functions (generators/sdf-fft/src/main/scala/WindowFunc.scala:41-65, SDFFFT.scala:53-66,233-234).  numpy restatement:
# integer arithmetic on the beats, so it stays exact.
WINDOW_NONE, WINDOW_HAMMING, WINDOW_HANNING, WINDOW_BLACKMAN, WINDOW_TRIANGULAR, WINDOW_USER = range(6)


def window_coeffs(kind, length, a0=0.0, a1=0.0, a2=0.0, dv=None):
    """the doubles of the window ROM: SDFFFT.scala:53-62 always calls the generators with their default isPeriodic = true"""
    if kind == WINDOW_NONE:
        return np.ones(length)
    if kind == WINDOW_USER:
        assert len(dv) == length, "Length of specified window function is not the same as fft size"
        return np.asarray(dv, dtype=np.float64)
    if length == 1:
        return np.ones(1)
    count = np.arange(length + 1, dtype=np.float64)
    if kind == WINDOW_HAMMING:          # hammingWindow(length, alpha, beta) (:43-47)
        w = a0 - a1 * np.cos(2 * np.pi * count / length)
    elif kind == WINDOW_HANNING:        # hanningWindow (:61-65)
        w = 0.5 * (1 - np.cos(2 * np.pi * count / length))
    elif kind == WINDOW_BLACKMAN:       # blackmanWindow (:55-59); its periodic branch has no .take(length): length + 1 values, the ROM is indexed 0..length-1
        w = a0 - a1 * np.cos(2 * np.pi * count / length) + a2 * np.cos(4 * np.pi * count / length)
    elif kind == WINDOW_TRIANGULAR:     # triangularWindow (:49-53)
        w = 1 - np.abs((count - length / 2.0) / (length / 2.0))
    else:
        raise ValueError(kind)
    return w[:length]


def window_raw(coeffs, width=16):
    """fromDoubleWithFixedWidth(t, FixedPoint(width, width - 2)): round half up, then the width-bit two's complement value"""
    raw = np.floor(np.ldexp(np.asarray(coeffs, dtype=np.float64), width - 2) + 0.5).astype(np.int64)
    return ((raw + (1 << (width - 1))) & ((1 << width) - 1)) - (1 << (width - 1))


def apply_window_words(beats, raw, width=16, data_width=16):
    """`in * window(cntWin)` on Cat(re16, im16) beats [rows, n]: full-precision product, the assignment to protoIQ floors away the
    width - 2 fractional bits and drops the MSBs (SDFFFT.scala:233-234); cntWin runs 0..n-1 over the first n ROM entries"""
    b = np.asarray(beats, dtype=np.uint32)
    n = b.shape[-1]
    w = np.asarray(raw, dtype=np.int64)[:n]
    half, mask = 1 << (data_width - 1), (1 << data_width) - 1
    def sx(v):
        return ((v.astype(np.int64) + half) & mask) - half
    re = sx((b >> 16) & mask); im = sx(b & mask)
    re = sx((re * w) >> (width - 2)); im = sx((im * w) >> (width - 2))
    return (((re & 0xFFFF) << 16) | (im & 0xFFFF)).astype(np.uint32)
