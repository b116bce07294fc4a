"""Host-side mirror of the reference's parameter case classes (same names, fields and defaults).

  FFTParams.fixed        generators/sdf-fft/src/main/scala/SDFFFTParams.scala:22-46,90-149
  MAGParams.fixed        generators/logMagMux/src/main/scala/LogMagMux.scala:14-49
  CFARParams             generators/cfar/src/main/scala/CFARParams.scala:10-49
  FixedPLFGParams        generators/plfg/src/main/scala/PLFGDspBlockMem.scala:20-72
  FixedNCOParams         generators/nco/src/main/scala/NCO.scala:14-83
  RspChainVanillaParameters / FftMagCfarVanillaParameters
                         src/main/scala/RspChain.scala:24-33, src/main/scala/FftMagCfarChain.scala:21-29

`require`s of the reference are enforced by the C library (RSP_EINVAL); the ones that only make sense on the Scala
side (array lengths) are checked here with the reference's messages."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

from . import _abi

# CFARType objects (CFARParams.scala:10-22)
CACFARType, GOSCFARType, GOSCACFARType = "CACFARType", "GOSCFARType", "GOSCACFARType"
# not a CFARType of the reference: the combined-window ordered statistic of its Scala golden cfarOS (CFARUtils.scala:159-196)
OSGoldenCFARType = "OSGoldenCFARType"
_CFAR_TYPES = {CACFARType: 0, GOSCFARType: 1, GOSCACFARType: 2, OSGoldenCFARType: 3}
# DecimType objects (SDFFFTParams.scala:18-20)
DIFDecimType, DITDecimType = "DIFDecimType", "DITDecimType"
# trim types
RoundHalfUp, Convergent, RoundDown = "RoundHalfUp", "Convergent", "RoundDown"


def log2Up(n: int) -> int:
    """chisel3.util.log2Up"""
    return max(1, (int(n) - 1).bit_length())


@dataclass(frozen=True)
class AddressSet:
    base: int
    mask: int

    def contains(self, addr: int) -> bool:
        return (addr ^ self.base) & ~self.mask == 0


@dataclass(frozen=True)
class FixedPoint:
    """chisel3.experimental.FixedPoint(width.W, binaryPoint.BP) as a type descriptor"""
    width: int
    binaryPoint: int


@dataclass
class WindowFunctionType:
    """generators/sdf-fft/src/main/scala/WindowFunc.scala:10-38 (`isPeriodic` is carried but, as in the reference's SDFFFT.scala:53-62,
    the ROM is always built from the periodic form)"""
    kind: int = 0
    dataWidth: int = 16
    a0: float = 0.0
    a1: float = 0.0
    a2: float = 0.0
    isPeriodic: bool = True
    dv: Optional[List[float]] = None

    def to_c(self):
        return (self.kind, self.dataWidth, float(self.a0), float(self.a1), float(self.a2))


class WindowFunctionTypes:
    @staticmethod
    def Hamming(dataWidth_tmp=16, alpha=0.54, beta=0.46, isPeriodic=True):
        return WindowFunctionType(1, dataWidth_tmp, alpha, beta, 0.0, isPeriodic)

    @staticmethod
    def Hanning(dataWidth_tmp=16, isPeriodic=True):
        return WindowFunctionType(2, dataWidth_tmp, 0.0, 0.0, 0.0, isPeriodic)

    @staticmethod
    def Blackman(dataWidth_tmp=16, a0=0.42, a1=0.5, a2=0.08, isPeriodic=True):
        return WindowFunctionType(3, dataWidth_tmp, a0, a1, a2, isPeriodic)

    @staticmethod
    def Triangular(dataWidth_tmp=16, isPeriodic=True):
        return WindowFunctionType(4, dataWidth_tmp, 0.0, 0.0, 0.0, isPeriodic)

    @staticmethod
    def User(dataWidth_tmp=16, dv=()):
        return WindowFunctionType(5, dataWidth_tmp, dv=list(dv))

    @staticmethod
    def NoWindow(dataWidth_tmp=16):
        return WindowFunctionType(0, dataWidth_tmp)


@dataclass
class FFTParams:
    numPoints: int = 2
    dataWidth: int = 16
    binPoint: int = 14
    twiddleWidth: int = 16
    keepMSBorLSBReg: bool = False
    keepMSBorLSB: List[bool] = field(default_factory=lambda: [True])
    overflowReg: bool = False
    fftType: str = "sdf"
    decimType: str = DIFDecimType
    sdfRadix: str = "2^2"
    runTimeR22: Optional[bool] = False
    expandLogic: List[int] = field(default_factory=lambda: [0])
    runTime: bool = False
    trimType: str = RoundHalfUp
    numAddPipes: int = 0
    numMulPipes: int = 0
    fftDir: bool = True
    fftDirReg: bool = False
    use4Muls: bool = False
    useBitReverse: bool = False
    minSRAMdepth: int = 0
    windowFunc: Optional["WindowFunctionType"] = None    # None = WindowFunctionTypes.None()
    arith: str = "fixed"       # "fixed" (FFTParams.fixed) or "fp32" (FFTParams.DSPReal run in single precision)
    windowRam: bool = False    # set by FFTBlockWithWindowing (chain.py): the block carries the run-time coefficient RAM + doWindowing register
    windowRamWidth: int = 16   # width of protoWin for that RAM (FixedPoint(w, w - 2)), WindowFunctionTypes.None(dataWidth_tmp)

    @staticmethod
    def fixed(**kw) -> "FFTParams":
        n = kw.get("numPoints", 2)
        kw.setdefault("keepMSBorLSB", [True] * log2Up(n))
        kw.setdefault("expandLogic", [0] * log2Up(n))
        p = FFTParams(**kw)
        p.checkNumPointsPow2()
        # SDFFFTParams.scala:53-86
        assert len(p.expandLogic) == log2Up(p.numPoints), "Size of expandLogic Array must be equal to log2 of numPoints"
        assert len(p.keepMSBorLSB) == log2Up(p.numPoints), "Size of keepMSBorLSB Array must be equal to log2 of numPoints"
        assert p.fftType == "sdf", "FFT type sdf is the only one supported"
        assert p.sdfRadix in ("2", "2^2"), "Radix must be 2 or 2^2"
        return p

    @staticmethod
    def DSPReal(**kw) -> "FFTParams":
        kw["arith"] = "fp32"
        return FFTParams.fixed(**kw)

    def checkNumPointsPow2(self):
        assert self.numPoints > 0 and self.numPoints & (self.numPoints - 1) == 0, "number of points must be a power of 2"

    def to_c(self) -> _abi.FftParams:
        mask = 0
        for i, k in enumerate(self.keepMSBorLSB):
            mask |= (1 if k else 0) << i
        return _abi.FftParams(
            self.numPoints, self.dataWidth, self.binPoint, self.twiddleWidth, _abi.TRIM[self.trimType],
            int(self.use4Muls), int(self.useBitReverse), int(self.fftDir), int(self.runTime), mask,
            0 if self.decimType == DIFDecimType else 1, 4 if self.sdfRadix == "2^2" else 2,
            sum((1 if e else 0) << i for i, e in enumerate(self.expandLogic)), int(self.keepMSBorLSBReg), int(self.fftDirReg),
            *(self.windowFunc.to_c() if self.windowFunc is not None else (0, self.windowRamWidth, 0.0, 0.0, 0.0)), int(self.overflowReg),
            int(self.windowRam))


@dataclass
class MAGParams:
    dataWidth: int = 16
    binPoint: int = 14
    dataWidthLog: int = 16
    binPointLog: int = 9
    log2LookUpWidth: int = 9
    useLast: bool = True
    numAddPipes: int = 1
    numMulPipes: int = 1

    @staticmethod
    def fixed(**kw) -> "MAGParams":
        return MAGParams(**kw)

    def to_c(self) -> _abi.MagParams:
        return _abi.MagParams(self.dataWidth, self.binPoint, self.dataWidthLog, self.binPointLog, self.log2LookUpWidth)


@dataclass
class CFARParams:
    protoIn: FixedPoint = FixedPoint(16, 0)
    protoThreshold: FixedPoint = FixedPoint(16, 0)
    protoScaler: FixedPoint = FixedPoint(16, 0)
    CFARAlgorithm: str = GOSCACFARType
    includeCASH: bool = False
    minSubWindowSize: Optional[int] = None
    leadLaggWindowSize: int = 16
    sendCut: bool = False
    guardWindowSize: int = 4
    fftSize: int = 1024
    retiming: bool = False
    numAddPipes: int = 0
    numMulPipes: int = 0

    def to_c(self) -> _abi.CfarParams:
        return _abi.CfarParams(
            self.protoIn.width, self.protoIn.binaryPoint, self.protoThreshold.width, self.protoThreshold.binaryPoint,
            self.protoScaler.width, self.protoScaler.binaryPoint, _CFAR_TYPES[self.CFARAlgorithm], int(self.includeCASH),
            int(self.minSubWindowSize or 0), self.leadLaggWindowSize, int(self.sendCut), self.guardWindowSize, self.fftSize)


@dataclass
class FixedPLFGParams:
    maxNumOfSegments: int = 4
    maxNumOfDifferentChirps: int = 8
    maxNumOfRepeatedChirps: int = 8
    maxChirpOrdinalNum: int = 4
    maxNumOfFrames: int = 4
    maxNumOfSamplesWidth: int = 8
    outputWidthInt: int = 16
    outputWidthFrac: int = 0

    def to_c(self) -> _abi.PlfgParams:
        return _abi.PlfgParams(self.maxNumOfSegments, self.maxNumOfDifferentChirps, self.maxNumOfRepeatedChirps,
                               self.maxChirpOrdinalNum, self.maxNumOfFrames, self.maxNumOfSamplesWidth,
                               self.outputWidthInt, self.outputWidthFrac)


@dataclass
class FixedNCOParams:
    tableSize: int = 128
    tableWidth: int = 16
    phaseWidth: int = 9
    rasterizedMode: bool = False
    nInterpolationTerms: int = 0
    ditherEnable: bool = False
    syncROMEnable: bool = False
    phaseAccEnable: bool = True
    roundingMode: str = RoundHalfUp
    pincType: str = "Streaming"
    poffType: str = "Fixed"
    useMultiplier: bool = False
    useQAM: bool = False

    def __post_init__(self):
        assert self.tableWidth >= 3                                     # NCO.scala:54
        if self.rasterizedMode:                                         # phaseW = log2Ceil(tableSize) + 2 (NCO.scala:56-62)
            self.phaseWidth = log2Up(self.tableSize) + 2 if self.tableSize > 1 else 2

    def to_c(self) -> _abi.NcoParams:
        if self.rasterizedMode:
            self.phaseWidth = log2Up(self.tableSize) + 2 if self.tableSize > 1 else 2
        return _abi.NcoParams(self.tableSize, self.tableWidth, self.phaseWidth, int(self.phaseAccEnable),
                              int(self.rasterizedMode), self.nInterpolationTerms, int(self.ditherEnable),
                              int(self.useMultiplier), int(self.useQAM), _abi.TRIM[self.roundingMode])


@dataclass
class FftMagCfarVanillaParameters:
    fftParams: FFTParams
    magParams: MAGParams
    cfarParams: CFARParams
    fftAddress: AddressSet = AddressSet(0x30000100, 0xFF)
    magAddress: AddressSet = AddressSet(0x30000200, 0xFF)
    cfarAddress: AddressSet = AddressSet(0x30002000, 0xFFF)
    beatBytes: int = 4


@dataclass
class RspChainVanillaParameters:
    plfgParams: FixedPLFGParams
    ncoParams: FixedNCOParams
    fftParams: FFTParams
    magParams: MAGParams
    cfarParams: CFARParams
    plfgAddress: AddressSet = AddressSet(0x30000000, 0xFF)
    plfgRAM: AddressSet = AddressSet(0x30001000, 0xFFF)
    ncoAddress: AddressSet = AddressSet(0x30000300, 0xF)
    fftAddress: AddressSet = AddressSet(0x30000100, 0xFF)
    magAddress: AddressSet = AddressSet(0x30000200, 0xFF)
    cfarAddress: AddressSet = AddressSet(0x30002000, 0xFFF)
    beatBytes: int = 4


def rsp_chain_vanilla_app_params() -> RspChainVanillaParameters:
    """RspChainVanillaApp (src/main/scala/RspChain.scala:80-147)"""
    return RspChainVanillaParameters(
        plfgParams=FixedPLFGParams(),
        ncoParams=FixedNCOParams(),
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=1024, useBitReverse=True, runTime=True,
                                  numAddPipes=1, numMulPipes=1, minSRAMdepth=1024, binPoint=0),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, 0), protoThreshold=FixedPoint(16, 0), protoScaler=FixedPoint(16, 0),
                              leadLaggWindowSize=64, guardWindowSize=4, fftSize=1024, minSubWindowSize=None,
                              includeCASH=False, CFARAlgorithm=CACFARType))


def fft_mag_cfar_chain_vanilla_app_params() -> FftMagCfarVanillaParameters:
    """FftMagCfarChainVanillaApp (src/main/scala/FftMagCfarChain.scala:76-120)"""
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=1024, useBitReverse=True, runTime=True,
                                  numAddPipes=1, numMulPipes=1, minSRAMdepth=1024, binPoint=12),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=12, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, 12), protoThreshold=FixedPoint(16, 12), protoScaler=FixedPoint(16, 12),
                              leadLaggWindowSize=64, guardWindowSize=4, sendCut=False, fftSize=1024,
                              minSubWindowSize=None, includeCASH=False, CFARAlgorithm=CACFARType))
