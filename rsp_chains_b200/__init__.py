"""rsp_chains_b200 -- the radar signal-processing hot path of rsp-chains (PLFG/NCO -> SDF-FFT -> LogMagMux -> CFAR) as
hand-written sm_100a CUDA behind a C ABI (include/rsp_chains.h), with a host-side mirror of the reference's parameter
case classes and chain/block interface.  CUDA-only: there is no CPU fallback."""
from . import _abi
from ._abi import RspError
from .chain import (FFTBlockWithWindowing, FftMagCfarChainVanilla, NcclComm, RspChainVanilla, gather_detections, gather_detections_nccl, host_alloc, host_free, shard_rows)
from .params import (AddressSet, CACFARType, CFARParams, Convergent, DIFDecimType, DITDecimType, FFTParams,
                     FftMagCfarVanillaParameters, FixedNCOParams, FixedPLFGParams, FixedPoint, GOSCACFARType, GOSCFARType, OSGoldenCFARType,
                     MAGParams, RoundDown, RoundHalfUp, RspChainVanillaParameters, WindowFunctionType, WindowFunctionTypes, fft_mag_cfar_chain_vanilla_app_params,
                     log2Up, rsp_chain_vanilla_app_params)
from .doppler import RangeDopplerChain
from .wire import BusOp, JtagReplay, RxFftMagCfarTxChain, beats_to_uart, jtag_decode, uart_to_beats

__all__ = [n for n in dir() if not n.startswith("_")]
