"""Host-side mirror of the reference's chain / block interface, on top of the C ABI.

  RspChainVanilla          src/main/scala/RspChain.scala:35-61        (plfg -> nco -> fft -> mag -> cfar)
  FftMagCfarChainVanilla   src/main/scala/FftMagCfarChain.scala:31-49 (fft -> mag -> cfar)
  AXI4FFTBlock / AXI4LogMagMuxBlock / AXI4CFARBlock / PLFGDspBlockMem + AXI4NCOLazyModuleBlock
                           the single DspBlocks (rsp_fft_run / rsp_mag_run / rsp_cfar_run / rsp_generate)

The memory-mapped register files are driven exactly as the reference testers do (memWriteWord(base + i*beatBytes, v),
src/test/scala/FftMagCfarChainTester.scala:82-132); the AXI4-Stream is a buffer of the same 32-bit beats.  torch is used
for device memory and streams only."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _abi
from .params import (AddressSet, CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedNCOParams, FixedPLFGParams,
                     MAGParams, RspChainVanillaParameters)


def _torch():
    import torch
    return torch


def _stream_ptr(stream) -> int:
    if stream is None:
        torch = _torch()
        return int(torch.cuda.current_stream().cuda_stream)
    return int(getattr(stream, "cuda_stream", stream))


class _Chain:
    """Common part of the two chain classes: handle, register bus, stream calls."""

    def __init__(self, c_params: _abi.ChainParams, address_map, beat_bytes: int, device: int = 0):
        self._lib = _abi.lib()
        self._h = C.c_void_p()
        self.device = int(device)
        self.beatBytes = beat_bytes
        self._address_map = address_map          # list of (AddressSet, block id or "plfg_ram")
        self._c_params = c_params
        _abi.check(self._lib.rsp_chain_create(C.byref(c_params), self.device, C.byref(self._h)))

    def _load_user_window(self, fft_params):
        """WindowFunctionTypes.User: the coefficients are part of the case class, the C ABI takes them right after creation"""
        w = fft_params.windowFunc
        if w is not None and w.kind == 5:
            assert len(w.dv) == fft_params.numPoints, "Length of specified window function is not the same as fft size"
            arr = (C.c_double * len(w.dv))(*w.dv)
            _abi.check(self._lib.rsp_fft_window_user(self._h, arr, len(w.dv)))

    # ---- life cycle
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.rsp_chain_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- register bus (AXI4MasterModel.memWriteWord / memReadWord)
    def _decode(self, addr: int):
        for aset, blk in self._address_map:
            if aset.contains(addr):
                off = addr - aset.base
                if off % self.beatBytes:
                    raise ValueError(f"unaligned register address 0x{addr:x}")
                return blk, off // self.beatBytes
        raise ValueError(f"address 0x{addr:x} decodes to no block")

    def memWriteWord(self, addr: int, value: int):
        blk, idx = self._decode(addr)
        if blk == "plfg_ram":
            _abi.check(self._lib.rsp_plfg_ram_write(self._h, idx, int(value) & 0xFFFFFFFF))
        else:
            _abi.check(self._lib.rsp_reg_write(self._h, blk, idx, int(value) & 0xFFFFFFFF))

    def memReadWord(self, addr: int) -> int:
        blk, idx = self._decode(addr)
        if blk in ("plfg_ram", "fft_win_ram"):
            raise ValueError("this RAM is write-only on this bus")
        v = C.c_uint32(0)
        _abi.check(self._lib.rsp_reg_read(self._h, blk, idx, C.byref(v)))
        return v.value

    CFAR_REGS = dict(fftWin=0, thresholdScaler=1, logOrLinearMode=2, divSum=3, peakGrouping=4, cfarAlgorithm=5, cfarMode=6,
                     windowCells=7, guardCells=8, indexLagg=9, indexLead=10, subWindowSize=11)

    def configure(self, log_n: int, sel: int, cfar_regs: dict):
        """write the run-time registers through the bus exactly as the reference testers do
        (src/test/scala/FftMagCfarChainTester.scala:82-132): fftSize, sel, then the CFAR register file by name"""
        p = self.params
        if p.fftParams.runTime:
            self.memWriteWord(p.fftAddress.base, log_n)
        self.memWriteWord(p.magAddress.base, sel)
        for k, v in cfar_regs.items():
            self.memWriteWord(p.cfarAddress.base + self.CFAR_REGS[k] * p.beatBytes, v)

    def reg_write(self, blk: int, idx: int, value: int):
        _abi.check(self._lib.rsp_reg_write(self._h, blk, idx, int(value) & 0xFFFFFFFF))

    def reg_read(self, blk: int, idx: int) -> int:
        v = C.c_uint32(0)
        _abi.check(self._lib.rsp_reg_read(self._h, blk, idx, C.byref(v)))
        return v.value

    def set_threshold_scaler_f32(self, scaler: float):
        bits = int(np.float32(scaler).view(np.uint32))
        _abi.check(self._lib.rsp_reg_write(self._h, _abi.BLK_CFAR, _abi.CFAR_REG_THRESHOLDSCALER_F32, bits))

    # ---- introspection
    @property
    def fixed(self) -> bool:
        return self._c_params.arith == _abi.ARITH_FIXED

    @property
    def fftSize(self) -> int:
        """run-time transform size N (the fftSize register holds log2 N)"""
        if self._c_params.fft.runTime:
            return 1 << self.reg_read(_abi.BLK_FFT, 0)
        return self._c_params.fft.numPoints

    @property
    def kernel_launches(self) -> int:
        return int(self._lib.rsp_kernel_launches(self._h))

    @property
    def variant(self) -> str:
        return self._lib.rsp_variant_name(self._h).decode()

    @property
    def bytes_per_chirp(self) -> int:
        return int(self._lib.rsp_bytes_per_chirp(self._h))

    # ---- stream: device buffers (torch tensors on self.device)
    def _dets_struct(self, det_list, det_count):
        if det_list is None:
            return None
        return _abi.Dets(det_list.data_ptr(), det_count.data_ptr(), det_list.shape[0])

    def run(self, d_in, n_chirps: int, n_rx: int = 1, out=None, peak_mask=None, det_list=None, det_count=None, stream=None):
        """fft -> mag -> cfar over n_rx * n_chirps chirps of fftSize beats (rsp_run).  All tensors live on the device.
        fixed: d_in int32/uint32 beats Cat(re16, im16); out int32 beats Cat(threshold, fftBin, peak).
        fp32 : d_in float32 [..., 2]; out float32 thresholds; peak_mask int32 words (1 bit per bin).
        det_list: int32 tensor [capacity, 4] (chirp, bin, threshold, cut); det_count: int32 tensor [1], zeroed by the caller."""
        d = self._dets_struct(det_list, det_count)
        _abi.check(self._lib.rsp_run(self._h, d_in.data_ptr(), n_chirps, n_rx, out.data_ptr() if out is not None else None,
                                     peak_mask.data_ptr() if peak_mask is not None else None,
                                     C.byref(d) if d is not None else None, _stream_ptr(stream)))
        return out

    def fft(self, d_in, n_chirps: int, out, stream=None):
        """AXI4FFTBlock alone (rsp_fft_run)"""
        _abi.check(self._lib.rsp_fft_run(self._h, d_in.data_ptr(), n_chirps, out.data_ptr(), _stream_ptr(stream)))
        return out

    @property
    def fft_in_beat_bytes(self) -> int:
        """bytes per input beat of the FFT block: 4 (Cat(re16, im16)), 8 for the wide block (dataWidth / twiddleWidth > 16, grow stages)"""
        return int(self._lib.rsp_fft_in_beat_bytes(self._h))

    @property
    def fft_out_beat_bytes(self) -> int:
        """bytes per output beat of the FFT block: 4, 8, or 16 when Cat(re, im) of the last stage's type needs more than 64 bits"""
        return int(self._lib.rsp_fft_out_beat_bytes(self._h))

    def fft_overflow_flags(self, stream=None) -> int:
        """per-stage overflow status of the most recent call (FFTParams.overflowReg; SDFChainRadix22.scala:473-486), all stages"""
        m = C.c_uint32(0)
        _abi.check(self._lib.rsp_fft_overflow_flags(self._h, C.byref(m), _stream_ptr(stream)))
        return int(m.value)

    def mag(self, d_in, n_beats: int, out, stream=None):
        """AXI4LogMagMuxBlock alone (rsp_mag_run)"""
        _abi.check(self._lib.rsp_mag_run(self._h, d_in.data_ptr(), n_beats, out.data_ptr(), _stream_ptr(stream)))
        return out

    def cfar(self, d_in, n_chirps: int, out=None, peak_mask=None, det_list=None, det_count=None, stream=None):
        """AXI4CFARBlock alone (rsp_cfar_run)"""
        d = self._dets_struct(det_list, det_count)
        _abi.check(self._lib.rsp_cfar_run(self._h, d_in.data_ptr(), n_chirps, out.data_ptr() if out is not None else None,
                                          peak_mask.data_ptr() if peak_mask is not None else None,
                                          C.byref(d) if d is not None else None, _stream_ptr(stream)))
        return out

    # ---- slow-time dimension (not in the reference)
    def corner_turn(self, d_in, n_rows: int, n_cols: int, out, batch: int = 1, stream=None):
        """out[b][c][r] = d_in[b][r][c] (rsp_corner_turn); element size from the tensor (2, 4 or 8 bytes; float32 [..., 2] counts as 8)"""
        eb = d_in.element_size() * (2 if (d_in.dim() and d_in.shape[-1] == 2 and d_in.is_floating_point()) else 1)
        _abi.check(self._lib.rsp_corner_turn(self._h, d_in.data_ptr(), out.data_ptr(), n_rows, n_cols, batch, eb, _stream_ptr(stream)))
        return out

    def mag_combine(self, d_in, n_beats: int, n_rx: int, out, stream=None):
        """LogMagMux of n_rx maps + their mean (rsp_mag_combine_run): non-coherent RX combining"""
        _abi.check(self._lib.rsp_mag_combine_run(self._h, d_in.data_ptr(), n_beats, n_rx, out.data_ptr(), _stream_ptr(stream)))
        return out

    # ---- stream: host buffers (numpy), the call a reference tester makes
    def run_host(self, h_in: np.ndarray, n_chirps: int, n_rx: int = 1, out: Optional[np.ndarray] = None,
                 peak_mask: Optional[np.ndarray] = None, max_dets: int = 0):
        """rsp_run_host: numpy in, numpy out; returns (out, peak_mask, detections[n,4])"""
        n = self.fftSize
        rows = n_chirps * n_rx
        h_in = np.ascontiguousarray(h_in)
        if out is None:
            out = np.empty((rows, n), dtype=np.uint32 if self.fixed else np.float32)
        if peak_mask is None and not self.fixed:
            peak_mask = np.empty((rows, (n + 31) // 32), dtype=np.uint32)
        dets = (_abi.Detection * max(1, max_dets))()
        n_dets = C.c_uint32(0)
        _abi.check(self._lib.rsp_run_host(
            self._h, h_in.ctypes.data, n_chirps, n_rx, out.ctypes.data,
            peak_mask.ctypes.data if peak_mask is not None else None,
            dets if max_dets else None, max_dets, C.byref(n_dets)))
        k = min(n_dets.value, max_dets)
        det_arr = np.frombuffer(dets, dtype=np.int32).reshape(-1, 4)[:k].copy() if max_dets else np.zeros((0, 4), np.int32)
        return out, peak_mask, det_arr, n_dets.value

    def run_host_raw(self, in_ptr: int, n_chirps: int, n_rx: int, out_ptr: int, mask_ptr: int = 0):
        """rsp_run_host on raw (e.g. pinned, rsp_host_alloc) pointers"""
        _abi.check(self._lib.rsp_run_host(self._h, in_ptr, n_chirps, n_rx, out_ptr or None, mask_ptr or None, None, 0, None))


    def run_host_dets_raw(self, in_ptr: int, n_chirps: int, n_rx: int, max_dets: int) -> int:
        """rsp_run_host with h_out = NULL: the detection list (kept in self.last_detections, a ctypes array reused between calls)
        is the only result copied back; returns the number of detections found (may exceed max_dets = overflow)"""
        if getattr(self, "_det_buf_cap", 0) < max_dets:
            self._det_buf = (_abi.Detection * max(1, max_dets))()
            self._det_buf_cap = max_dets
        n_dets = C.c_uint32(0)
        _abi.check(self._lib.rsp_run_host(self._h, in_ptr, n_chirps, n_rx, None, None, self._det_buf, max_dets, C.byref(n_dets)))
        self.last_detections = self._det_buf
        return int(n_dets.value)


def host_alloc(nbytes: int, dtype=np.uint32) -> np.ndarray:
    """pinned host memory from rsp_host_alloc as a numpy array (freed with host_free)"""
    p = C.c_void_p()
    _abi.check(_abi.lib().rsp_host_alloc(C.byref(p), nbytes))
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype)
    arr.flags.writeable = True
    return arr


def host_free(arr: np.ndarray):
    _abi.lib().rsp_host_free(C.c_void_p(arr.ctypes.data))


def _c_params(fft: FFTParams, mag: MAGParams, cfar: CFARParams, plfg: Optional[FixedPLFGParams] = None,
              nco: Optional[FixedNCOParams] = None) -> _abi.ChainParams:
    p = _abi.ChainParams()
    p.arith = _abi.ARITH_FIXED if fft.arith == "fixed" else _abi.ARITH_FP32
    p.hasGenerator = int(plfg is not None)
    p.plfg = (plfg or FixedPLFGParams()).to_c()
    p.nco = (nco or FixedNCOParams()).to_c()
    p.fft = fft.to_c()
    p.mag = mag.to_c()
    p.cfar = cfar.to_c()
    return p


class FftMagCfarChainVanilla(_Chain):
    """fft -> mag -> cfar (src/main/scala/FftMagCfarChain.scala:31-49)"""

    def __init__(self, params: FftMagCfarVanillaParameters, device: int = 0):
        self.params = params
        amap = [(params.fftAddress, _abi.BLK_FFT), (params.magAddress, _abi.BLK_MAG), (params.cfarAddress, _abi.BLK_CFAR)]
        super().__init__(_c_params(params.fftParams, params.magParams, params.cfarParams), amap, params.beatBytes, device)
        self._load_user_window(params.fftParams)


class FFTBlockWithWindowing(_Chain):
    """The FFT block with a run-time window RAM (generators/sdf-fft/src/main/scala/FFTBlockWithWindowing.scala:18-176): same constructor
    arguments as the reference (csrAddress, ramAddress, params, beatBytes).  Registers: fftSize 0, fftDir 1, doWindowing 2,
    keepMSBorLSBReg 3, busy 4, overflowReg 5 (:134-152); memWriteWord(ramAddress.base + i * beatBytes, raw) writes coefficient i."""

    def __init__(self, csrAddress: AddressSet, ramAddress: AddressSet, params: FFTParams, beatBytes: int = 4, device: int = 0,
                 windowDataWidth: int = 16):
        import copy
        assert params.windowFunc is None or params.windowFunc.kind == 0, \
            "Parameter for window function inside SDFFFT module must be set to None type"
        fp = copy.copy(params)
        fp.windowRam = True
        fp.windowRamWidth = params.windowFunc.dataWidth if params.windowFunc is not None else windowDataWidth
        self.fftParams = fp
        self.csrAddress, self.ramAddress = csrAddress, ramAddress
        n = max(params.numPoints, 64)
        mag = MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9)
        cfar = CFARParams(leadLaggWindowSize=16, guardWindowSize=2, fftSize=n)
        amap = [(ramAddress, "fft_win_ram"), (csrAddress, _abi.BLK_FFT)]
        super().__init__(_c_params(fp, mag, cfar), amap, beatBytes, device)

    def memWriteWord(self, addr: int, value: int):
        blk, idx = self._decode(addr)
        if blk == "fft_win_ram":
            _abi.check(self._lib.rsp_fft_window_ram_write(self._h, idx, int(value) & 0xFFFFFFFF))
        else:
            _abi.check(self._lib.rsp_reg_write(self._h, blk, idx, int(value) & 0xFFFFFFFF))


class RspChainVanilla(_Chain):
    """plfg -> nco -> fft -> mag -> cfar (src/main/scala/RspChain.scala:35-61)"""

    def __init__(self, params: RspChainVanillaParameters, device: int = 0):
        self.params = params
        amap = [(params.plfgRAM, "plfg_ram"), (params.plfgAddress, _abi.BLK_PLFG), (params.ncoAddress, _abi.BLK_NCO),
                (params.fftAddress, _abi.BLK_FFT), (params.magAddress, _abi.BLK_MAG), (params.cfarAddress, _abi.BLK_CFAR)]
        super().__init__(_c_params(params.fftParams, params.magParams, params.cfarParams, params.plfgParams, params.ncoParams),
                         amap, params.beatBytes, device)
        self._load_user_window(params.fftParams)

    def generate(self, n_samples: int, nco_out=None, plfg_out=None, stream=None):
        """PLFGDspBlockMem -> AXI4NCOLazyModuleBlock alone (rsp_generate)"""
        _abi.check(self._lib.rsp_generate(self._h, n_samples, nco_out.data_ptr() if nco_out is not None else None,
                                          plfg_out.data_ptr() if plfg_out is not None else None, _stream_ptr(stream)))
        return nco_out, plfg_out

    def run_from_generator(self, n_chirps: int, n_rx: int = 1, out=None, peak_mask=None, det_list=None, det_count=None,
                           nco_out=None, stream=None):
        """the whole chain with the PLFG enabled (rsp_run_from_generator)"""
        d = self._dets_struct(det_list, det_count)
        _abi.check(self._lib.rsp_run_from_generator(
            self._h, n_chirps, n_rx, out.data_ptr() if out is not None else None,
            peak_mask.data_ptr() if peak_mask is not None else None, C.byref(d) if d is not None else None,
            nco_out.data_ptr() if nco_out is not None else None, _stream_ptr(stream)))
        return out

    def run_from_generator_static(self, n_chirps: int, n_rx: int = 1, out=None, peak_mask=None, det_list=None, det_count=None, stream=None):
        """rsp_run_from_generator_static: the same results with the scratch beats in a fixed device array of the library -- no cudaMalloc in
        the call (at most RSP_STATIC_SCRATCH_BEATS = 2^20 samples per call; calls that overlap on the GPU race on the array)"""
        d = self._dets_struct(det_list, det_count)
        _abi.check(self._lib.rsp_run_from_generator_static(
            self._h, n_chirps, n_rx, out.data_ptr() if out is not None else None,
            peak_mask.data_ptr() if peak_mask is not None else None, C.byref(d) if d is not None else None, _stream_ptr(stream)))
        return out


def gather_detections(det_list, det_count, group=None):
    """Gather the per-rank detection lists on every rank (torch.distributed all_gather; NCCL on GPUs, gloo on CPU).
    det_list: int32 [capacity, 4]; det_count: int32 [1].  Returns an int32 tensor [total, 5] = (rank, chirp, bin, thr, cut).
    This is the only inter-GPU exchange of the path: frames are sharded, detections are tiny."""
    torch = _torch()
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_local = int(min(int(det_count.item()), det_list.shape[0]))
    if world == 1:
        loc = det_list[:n_local]
        return torch.cat([torch.full((n_local, 1), rank, dtype=loc.dtype, device=loc.device), loc], dim=1)
    counts = [torch.zeros(1, dtype=torch.int64, device=det_list.device) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([n_local], dtype=torch.int64, device=det_list.device), group=group)
    n_max = max(int(c.item()) for c in counts)
    padded = torch.zeros((max(n_max, 1), 4), dtype=det_list.dtype, device=det_list.device)
    padded[:n_local] = det_list[:n_local]
    bufs = [torch.zeros_like(padded) for _ in range(world)]
    dist.all_gather(bufs, padded, group=group)
    parts = []
    for r, (b, c) in enumerate(zip(bufs, counts)):
        k = int(c.item())
        parts.append(torch.cat([torch.full((k, 1), r, dtype=b.dtype, device=b.device), b[:k]], dim=1))
    return torch.cat(parts, dim=0) if parts else torch.zeros((0, 5), dtype=det_list.dtype, device=det_list.device)


class NcclComm:
    """An ncclComm_t created through the C ABI (rsp_nccl_get_unique_id / rsp_nccl_comm_init_rank): what a host without its own NCCL
    binding (the Scala side of INTEGRATION.md) does.  The 128-byte unique id of rank 0 travels through torch.distributed here."""

    def __init__(self, device: int, group=None):
        torch = _torch()
        import torch.distributed as dist
        self._lib = _abi.lib()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        buf = C.create_string_buffer(128)
        if self.rank == 0:
            _abi.check(self._lib.rsp_nccl_get_unique_id(buf))
        if self.world > 1:
            box = [bytes(buf.raw)]
            dist.broadcast_object_list(box, src=0, group=group)
            buf = C.create_string_buffer(box[0], 128)
        self._h = C.c_void_p()
        _abi.check(self._lib.rsp_nccl_comm_init_rank(C.byref(self._h), self.world, buf, self.rank, int(device)))
        torch.cuda.synchronize()

    def close(self):
        if self._h and self._h.value:
            self._lib.rsp_nccl_comm_destroy(self._h)
            self._h = C.c_void_p()


def gather_detections_nccl(chain, comm: NcclComm, det_list, det_count, slot_capacity: int, stream=None):
    """rsp_gather_detections: the detection lists of all ranks through two ncclAllGather calls on `stream` (asynchronous).
    Returns (gathered int32 [world, slot_capacity, 4], counts int32 [world]) device tensors: rank r's detections are
    gathered[r, :min(counts[r], slot_capacity)]."""
    torch = _torch()
    gathered = torch.empty((comm.world, slot_capacity, 4), dtype=torch.int32, device=det_list.device)
    counts = torch.zeros(comm.world, dtype=torch.int32, device=det_list.device)
    d = _abi.Dets(det_list.data_ptr(), det_count.data_ptr(), det_list.shape[0])
    _abi.check(_abi.lib().rsp_gather_detections(chain._h, comm._h, C.byref(d), slot_capacity, gathered.data_ptr(), counts.data_ptr(),
                                                _stream_ptr(stream)))
    return gathered, counts


def shard_rows(n_units: int, world_size: int, rank: int):
    """contiguous block of units (frames / chirps) owned by `rank`: unit u goes to gpu u * world / n_units"""
    lo = (rank * n_units + world_size - 1) // world_size
    hi = ((rank + 1) * n_units + world_size - 1) // world_size
    return lo, hi
