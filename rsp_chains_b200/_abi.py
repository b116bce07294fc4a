"""ctypes binding of librsp_b200.so (include/rsp_chains.h).  Plain pointers and sizes only; torch is used by callers for
device memory and streams, never by this module.  There is no CPU path: if the library is missing, import fails loudly."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RSP_B200_LIB: another build of the library (A/B studies of compile-time kernel knobs, scripts/build_variant.sh)
LIB_PATH = os.environ.get("RSP_B200_LIB") or os.path.join(_HERE, "librsp_b200.so")

RSP_OK, RSP_EINVAL, RSP_ECUDA, RSP_ENOMEM, RSP_EUNSUPPORTED = 0, 1, 2, 3, 4
BLK_PLFG, BLK_NCO, BLK_FFT, BLK_MAG, BLK_CFAR = 0, 1, 2, 3, 4
ARITH_FIXED, ARITH_FP32 = 0, 1
CFAR_REG_THRESHOLDSCALER_F32 = 0x100

TRIM = {"RoundHalfUp": 0, "RoundDown": 1, "Floor": 1, "Convergent": 2, "RoundHalfToEven": 2, "RoundUp": 3,
        "RoundTowardsZero": 4, "RoundTowardsInfinity": 5, "RoundHalfDown": 6, "RoundHalfTowardsZero": 7,
        "RoundHalfTowardsInfinity": 8, "RoundHalfToOdd": 9, "NoTrim": 10}

EXPORTS = [
    "rsp_chain_params_default", "rsp_chain_create", "rsp_chain_destroy", "rsp_reg_write", "rsp_reg_read",
    "rsp_plfg_ram_write", "rsp_fft_window_user", "rsp_fft_window_ram_write", "rsp_run", "rsp_run_from_generator", "rsp_fft_run", "rsp_mag_run", "rsp_cfar_run",
    "rsp_generate", "rsp_run_host", "rsp_host_alloc", "rsp_host_free", "rsp_status_str", "rsp_last_error",
    "rsp_abi_version", "rsp_kernel_launches", "rsp_variant_name", "rsp_bytes_per_chirp",
    "rsp_fft_in_beat_bytes", "rsp_fft_out_beat_bytes", "rsp_fft_overflow_flags",
    "rsp_gather_detections", "rsp_nccl_get_unique_id", "rsp_nccl_comm_init_rank", "rsp_nccl_comm_destroy",
    "rsp_jtag_decode", "rsp_uart_to_beats", "rsp_beats_to_uart", "rsp_corner_turn", "rsp_mag_combine_run", "rsp_run_from_generator_static",
]


def _i32s(*names):
    return [(n, C.c_int32) for n in names]


class FftParams(C.Structure):
    _fields_ = _i32s("numPoints", "dataWidth", "binPoint", "twiddleWidth", "trimType", "use4Muls", "useBitReverse",
                     "fftDir", "runTime") + [("keepMSBorLSB", C.c_uint32)] + _i32s(
        "decimType", "sdfRadix") + [("expandLogic", C.c_uint32)] + _i32s("keepMSBorLSBReg", "fftDirReg", "windowFunc", "windowWidth") + [
        ("windowA0", C.c_double), ("windowA1", C.c_double), ("windowA2", C.c_double)] + _i32s("overflowReg", "windowRam")


class BusOp(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("addr", C.c_uint32), ("data_lo", C.c_uint32), ("data_hi", C.c_uint32)]


class MagParams(C.Structure):
    _fields_ = _i32s("dataWidth", "binPoint", "dataWidthLog", "binPointLog", "log2LookUpWidth")


class CfarParams(C.Structure):
    _fields_ = _i32s("inWidth", "inBinPoint", "thresholdWidth", "thresholdBinPoint", "scalerWidth", "scalerBinPoint",
                     "CFARAlgorithm", "includeCASH", "minSubWindowSize", "leadLaggWindowSize", "sendCut",
                     "guardWindowSize", "fftSize")


class PlfgParams(C.Structure):
    _fields_ = _i32s("maxNumOfSegments", "maxNumOfDifferentChirps", "maxNumOfRepeatedChirps", "maxChirpOrdinalNum",
                     "maxNumOfFrames", "maxNumOfSamplesWidth", "outputWidthInt", "outputWidthFrac")


class NcoParams(C.Structure):
    _fields_ = _i32s("tableSize", "tableWidth", "phaseWidth", "phaseAccEnable", "rasterizedMode",
                     "nInterpolationTerms", "ditherEnable", "useMultiplier", "useQAM", "roundingMode")


class ChainParams(C.Structure):
    _fields_ = _i32s("arith", "hasGenerator") + [("plfg", PlfgParams), ("nco", NcoParams), ("fft", FftParams),
                                                  ("mag", MagParams), ("cfar", CfarParams)]


class Detection(C.Structure):
    _fields_ = [("chirp", C.c_uint32), ("bin", C.c_uint32), ("threshold", C.c_int32), ("cut", C.c_int32)]


class Dets(C.Structure):
    _fields_ = [("list", C.c_void_p), ("count", C.c_void_p), ("capacity", C.c_uint32)]


class RspError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{status_name(status)}: {message}")
        self.status = status
        self.message = message


_lib = None


def lib():
    """Load the C-ABI library.  Raises if it is missing: the product has no CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(librsp_b200 is CUDA-only and has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        L.rsp_chain_params_default.argtypes = [C.POINTER(ChainParams)]
        L.rsp_chain_params_default.restype = None
        L.rsp_chain_create.argtypes = [C.POINTER(ChainParams), i32, C.POINTER(vp)]
        L.rsp_chain_destroy.argtypes = [vp]
        L.rsp_chain_destroy.restype = None
        L.rsp_reg_write.argtypes = [vp, i32, u32, u32]
        L.rsp_reg_read.argtypes = [vp, i32, u32, C.POINTER(u32)]
        L.rsp_plfg_ram_write.argtypes = [vp, u32, u32]
        L.rsp_fft_window_user.argtypes = [vp, C.POINTER(C.c_double), u32]
        L.rsp_fft_window_ram_write.argtypes = [vp, u32, u32]
        L.rsp_run.argtypes = [vp, vp, u32, u32, vp, vp, C.POINTER(Dets), vp]
        L.rsp_run_from_generator.argtypes = [vp, u32, u32, vp, vp, C.POINTER(Dets), vp, vp]
        L.rsp_fft_run.argtypes = [vp, vp, u32, vp, vp]
        L.rsp_mag_run.argtypes = [vp, vp, u64, vp, vp]
        L.rsp_cfar_run.argtypes = [vp, vp, u32, vp, vp, C.POINTER(Dets), vp]
        L.rsp_generate.argtypes = [vp, u64, vp, vp, vp]
        L.rsp_run_host.argtypes = [vp, vp, u32, u32, vp, vp, C.POINTER(Detection), u32, C.POINTER(u32)]
        L.rsp_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        L.rsp_host_free.argtypes = [vp]
        L.rsp_host_free.restype = None
        L.rsp_status_str.argtypes = [i32]
        L.rsp_status_str.restype = C.c_char_p
        L.rsp_last_error.restype = C.c_char_p
        L.rsp_abi_version.restype = i32
        L.rsp_kernel_launches.argtypes = [vp]
        L.rsp_kernel_launches.restype = u64
        L.rsp_variant_name.argtypes = [vp]
        L.rsp_variant_name.restype = C.c_char_p
        L.rsp_bytes_per_chirp.argtypes = [vp]
        L.rsp_bytes_per_chirp.restype = u64
        L.rsp_fft_in_beat_bytes.argtypes = [vp]
        L.rsp_fft_in_beat_bytes.restype = u32
        L.rsp_fft_out_beat_bytes.argtypes = [vp]
        L.rsp_fft_out_beat_bytes.restype = u32
        L.rsp_fft_overflow_flags.argtypes = [vp, C.POINTER(u32), vp]
        L.rsp_gather_detections.argtypes = [vp, vp, C.POINTER(Dets), u32, vp, vp, vp]
        L.rsp_nccl_get_unique_id.argtypes = [C.c_char_p]
        L.rsp_nccl_comm_init_rank.argtypes = [C.POINTER(vp), i32, C.c_char_p, i32, i32]
        L.rsp_nccl_comm_destroy.argtypes = [vp]
        L.rsp_nccl_comm_destroy.restype = None
        L.rsp_jtag_decode.argtypes = [C.c_char_p, C.c_size_t, C.c_uint8, C.c_uint8, C.c_uint8, u32, u32, u32, C.POINTER(BusOp), u32, C.POINTER(u32)]
        L.rsp_uart_to_beats.argtypes = [C.c_char_p, C.c_size_t, u32, i32, vp, C.POINTER(C.c_size_t)]
        L.rsp_beats_to_uart.argtypes = [vp, C.c_size_t, u32, i32, vp]
        L.rsp_corner_turn.argtypes = [vp, vp, vp, u32, u32, u32, u32, vp]
        L.rsp_mag_combine_run.argtypes = [vp, vp, u64, u32, vp, vp]
        L.rsp_run_from_generator_static.argtypes = [vp, u32, u32, vp, vp, C.POINTER(Dets), vp]
        _lib = L
    return _lib


def status_name(st):
    return lib().rsp_status_str(int(st)).decode()


def check(st):
    if st != RSP_OK:
        raise RspError(st, lib().rsp_last_error().decode())
