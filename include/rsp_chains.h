/*
 * rsp_chains.h -- C ABI of librsp_b200.so, the B200 (sm_100a) implementation of the rsp-chains hot path
 *                 PLFG -> NCO -> SDF-FFT -> LogMagMux -> CFAR, batched over chirps / RX channels / frames.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ / torch types.  Each entry point names the
 * reference interface it replaces (file:line relative to the rsp-chains repository).  The reference is a
 * Chisel hardware generator: a "block" there is a rocket-chip DspBlock with one AXI4-Stream node (32-bit
 * beats + `last`) and one memory-mapped register file; a "chain" is blocks bound with `:=`.  Here a chain is
 * an opaque handle, the compile-time case classes are the structs below, the register files are
 * rsp_reg_write/rsp_reg_read with the reference's word offsets, and the stream is a device (or host) buffer
 * of the very same 32-bit beats, one chirp after another.
 *
 * Conventions
 *   - every function returns rsp_status (0 = OK); rsp_last_error() gives a message for the calling thread
 *   - the caller owns every buffer it passes; the handle owns tables, registers and scratch
 *   - device entry points are asynchronous on the given stream (a cudaStream_t passed as void*; NULL = the
 *     legacy default stream); host entry points (rsp_*_host) return after the results are in the host buffer
 *   - a handle serves ONE thread and ONE stream at a time: its scratch (generator beats, segment list, detection staging,
 *     the chunk buffers of rsp_run_host) is shared by every call, so two calls of one handle on different streams must be
 *     ordered by the caller; distinct handles are independent.  A buffer that has to grow is reallocated inside the call
 *     after a cudaDeviceSynchronize (first call, or a larger batch than any before)
 *   - every entry point runs on the handle's device and restores the caller's current device before it returns
 *   - rsp_last_error() points into thread-local storage: copy the text before the thread ends
 *   - parameter checks reproduce the reference's `require`s / `assert`s and return RSP_EINVAL
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with RSP_ECUDA
 */
#ifndef RSP_CHAINS_H
#define RSP_CHAINS_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RSP_ABI_VERSION 3

typedef enum {
  RSP_OK = 0,
  RSP_EINVAL = 1,       /* a reference `require`/`assert` would have fired, or a bad argument */
  RSP_ECUDA = 2,        /* CUDA runtime error (no device, launch failure, ...) */
  RSP_ENOMEM = 3,
  RSP_EUNSUPPORTED = 4  /* legal in the reference generator, not built here (listed in DESIGN.md) */
} rsp_status;

/* blocks of the chain, in composition order (src/main/scala/RspChain.scala:41-58) */
typedef enum { RSP_BLK_PLFG = 0, RSP_BLK_NCO = 1, RSP_BLK_FFT = 2, RSP_BLK_MAG = 3, RSP_BLK_CFAR = 4 } rsp_block;

/* arithmetic of the data path: the FixedPoint generator, or the "DspReal golden" flavour run in fp32 */
typedef enum { RSP_ARITH_FIXED = 0, RSP_ARITH_FP32 = 1 } rsp_arith;

/* dsptools TrimType (generators/sdf-fft/src/test/scala/Radix22SQNRSpec.scala:184) */
typedef enum {
  RSP_TRIM_ROUND_HALF_UP = 0, RSP_TRIM_FLOOR = 1 /* RoundDown */, RSP_TRIM_CONVERGENT = 2 /* RoundHalfToEven */,
  RSP_TRIM_ROUND_UP = 3, RSP_TRIM_ROUND_TOWARDS_ZERO = 4, RSP_TRIM_ROUND_TOWARDS_INFINITY = 5,
  RSP_TRIM_ROUND_HALF_DOWN = 6, RSP_TRIM_ROUND_HALF_TOWARDS_ZERO = 7, RSP_TRIM_ROUND_HALF_TOWARDS_INFINITY = 8,
  RSP_TRIM_ROUND_HALF_TO_ODD = 9, RSP_TRIM_NO_TRIM = 10
} rsp_trim;

/* ---- compile-time tier: C mirrors of the case classes ------------------------------------------------ */

/* FFTParams.fixed (generators/sdf-fft/src/main/scala/SDFFFTParams.scala:22-46,90-149) */
typedef struct {
  int32_t numPoints;        /* power of two, 2..16384                                                    */
  int32_t dataWidth;        /* 2..16 in a chain; 17..32 for the FFT block alone (wide beats, see rsp_fft_run) */
  int32_t binPoint;
  int32_t twiddleWidth;     /* 3..16 in a chain, up to 32 for the wide FFT block; twiddles are Q2.(twiddleWidth-2) (:116) */
  int32_t trimType;         /* rsp_trim, used by div2 / context_* / trimBinary                            */
  int32_t use4Muls;         /* complexUse4Muls                                                            */
  int32_t useBitReverse;    /* 1: natural-order output (BitReversePingPong), 0: bit-reversed              */
  int32_t fftDir;           /* 1 = FFT, 0 = IFFT (reset value of the fftDir register)                     */
  int32_t runTime;          /* 1: the fftSize register may select a smaller transform                     */
  uint32_t keepMSBorLSB;    /* bit i = keepMSBorLSB(i): 1 scale by 1/2 with trimType, 0 keep LSBs (wrap)   */
  int32_t decimType;        /* 0 = DIFDecimType, 1 = DITDecimType (either radix, fixed point and fp32)                   */
  int32_t sdfRadix;         /* 4 = "2^2" (SDFChainRadix22.scala), 2 = "2" (SDFChainRadix2.scala)            */
  uint32_t expandLogic;     /* bit i = expandLogic(i): stage i keeps the grown butterfly outputs in a type one bit wider
                             * (SDFFFTParams.scala:119-123, SDFChainRadix22.scala:453-457); radix 2^2 DIF, FFT block alone (wide beats) */
  int32_t keepMSBorLSBReg;  /* 1: the keepMSBorLSBReg register overrides keepMSBorLSB (FFTBlock.scala:92-93)   */
  int32_t fftDirReg;        /* 1: the fftDir register overrides fftDir (FFTBlock.scala:90-91)                 */
  /* windowFunc (generators/sdf-fft/src/main/scala/WindowFunc.scala:10-38): the coefficient ROM of numPoints entries of type
   * protoWin = FixedPoint(windowWidth, windowWidth - 2) (SDFFFTParams.scala:118; doubles in the fp32 flavour), multiplied into
   * the forward transform's input, `in * window(cntWin)` truncated on assignment (SDFFFT.scala:53-66,233-234).                */
  int32_t windowFunc;       /* rsp_window; RSP_WINDOW_NONE = no multiply                                    */
  int32_t windowWidth;      /* WindowFunctionType.dataWidth, 3..16                                           */
  double windowA0, windowA1, windowA2;   /* Hamming: alpha, beta; Blackman: a0, a1, a2 (WindowFunc.scala:13-28)      */
  int32_t overflowReg;      /* FFTParams.overflowReg: the per-stage overflow status is computed (RSP_FFT_REG_OVERFLOW, rsp_fft_overflow_flags);
                             * such a chain runs on the run-time-parameter kernels                                                      */
  int32_t windowRam;        /* 1: the block is FFTBlockWithWindowing (generators/sdf-fft/src/main/scala/FFTBlockWithWindowing.scala:18-176): a RAM of
                             * numPoints protoWin coefficients written at run time (rsp_fft_window_ram_write) multiplies the input,
                             * `inComplex.real * winCoeff` (:159-166), when the doWindowing register is 1 -- in both transform directions;
                             * requires windowFunc = RSP_WINDOW_NONE (:19); the register file is RSP_FFTW_REG_* below                  */
} rsp_fft_params;
typedef enum { RSP_WINDOW_NONE = 0, RSP_WINDOW_HAMMING = 1, RSP_WINDOW_HANNING = 2, RSP_WINDOW_BLACKMAN = 3, RSP_WINDOW_TRIANGULAR = 4,
               RSP_WINDOW_USER = 5 } rsp_window;

/* MAGParams.fixed (generators/logMagMux/src/main/scala/LogMagMux.scala:14-49) */
typedef struct {
  int32_t dataWidth, binPoint;        /* proto          */
  int32_t dataWidthLog, binPointLog;  /* protoLog       */
  int32_t log2LookUpWidth;
} rsp_mag_params;

/* CFARParams (generators/cfar/src/main/scala/CFARParams.scala:24-49) */
typedef enum { RSP_CFAR_CA = 0 /* CACFARType */, RSP_CFAR_GOS = 1 /* GOSCFARType */, RSP_CFAR_GOSCA = 2 /* GOSCACFARType */,
               /* not a CFARType of the reference: the combined-window ordered statistic of its Scala golden model cfarOS
                * (generators/cfar/src/main/scala/CFARUtils.scala:159-196): threshold = scaler (*|+) the cell at ascending index
                * floor(1.5 * windowCells) of the 2 * windowCells reference cells (lead ++ lagg), 0 and no peak within
                * windowCells + guardCells of either end.  Registers as for the Lis core; cfarMode / indexLagg / indexLead unused. */
               RSP_CFAR_OS_GOLDEN = 3 } rsp_cfar_type;
typedef struct {
  int32_t inWidth, inBinPoint;               /* protoIn         */
  int32_t thresholdWidth, thresholdBinPoint; /* protoThreshold  */
  int32_t scalerWidth, scalerBinPoint;       /* protoScaler     */
  int32_t CFARAlgorithm;                     /* rsp_cfar_type   */
  int32_t includeCASH;
  int32_t minSubWindowSize;                  /* 0 = None        */
  int32_t leadLaggWindowSize;                /* power of two    */
  int32_t sendCut;
  int32_t guardWindowSize;
  int32_t fftSize;                           /* power of two; width of the fftBin field of the output beat */
} rsp_cfar_params;

/* FixedPLFGParams (generators/plfg/src/main/scala/PLFGDspBlockMem.scala:20-72) */
typedef struct {
  int32_t maxNumOfSegments, maxNumOfDifferentChirps, maxNumOfRepeatedChirps, maxChirpOrdinalNum, maxNumOfFrames;
  int32_t maxNumOfSamplesWidth, outputWidthInt, outputWidthFrac;
} rsp_plfg_params;

/* FixedNCOParams, standard-mode or rasterized table + accumulator (generators/nco/src/main/scala/NCO.scala:14-83) */
typedef struct {
  int32_t tableSize, tableWidth, phaseWidth;
  int32_t phaseAccEnable;
  int32_t rasterizedMode;  /* NCOTableRasterizedMode (NCOTable.scala:322-464): tableSize need not be a power of two, address = phase mod
                            * tableSize, quadrant = (phase / tableSize) mod 4; phaseWidth must be log2Ceil(tableSize) + 2 (NCO.scala:56-62)      */
  int32_t nInterpolationTerms, ditherEnable;   /* non-zero -> RSP_EUNSUPPORTED */
  int32_t useMultiplier;   /* the output multiplier and its two registers RSP_NCO_REG_* (NCOLazyModuleBlock.scala:262-269,354-364)   */
  int32_t useQAM;          /* non-zero -> RSP_EUNSUPPORTED */
  int32_t roundingMode;    /* rsp_trim, FixedNCOParams.roundingMode (default RoundHalfUp = 0): trim of the multiplier's product     */
} rsp_nco_params;

/* RspChainVanillaParameters / FftMagCfarVanillaParameters (src/main/scala/RspChain.scala:24-33,
 * src/main/scala/FftMagCfarChain.scala:21-29) without the AddressSets (there is no bus) */
typedef struct {
  int32_t arith;          /* rsp_arith */
  int32_t hasGenerator;   /* 1: RspChainVanilla (plfg+nco in front), 0: FftMagCfarChainVanilla */
  rsp_plfg_params plfg;
  rsp_nco_params nco;
  rsp_fft_params fft;
  rsp_mag_params mag;
  rsp_cfar_params cfar;
} rsp_chain_params;

/* defaults = RspChainVanillaApp (src/main/scala/RspChain.scala:80-147): 1024 points, 16-bit, binPoint 0 */
void rsp_chain_params_default(rsp_chain_params *p);

/* ---- run-time tier: register word indices (= byte offset / beatBytes in the reference regmaps) -------- */
enum { /* AXI4FFTBlock (generators/sdf-fft/src/main/scala/FFTBlock.scala:86-102) */
  RSP_FFT_REG_FFTSIZE = 0 /* log2 of the run-time size */, RSP_FFT_REG_FFTDIR = 1, RSP_FFT_REG_KEEPMSBORLSB = 2,
  RSP_FFT_REG_BUSY = 3 /* RO */, RSP_FFT_REG_OVERFLOW = 4 /* RO */ };
enum { /* FFTBlockWithWindowing (generators/sdf-fft/src/main/scala/FFTBlockWithWindowing.scala:134-152): used instead when windowRam = 1 */
  RSP_FFTW_REG_FFTSIZE = 0, RSP_FFTW_REG_FFTDIR = 1, RSP_FFTW_REG_DOWINDOWING = 2, RSP_FFTW_REG_KEEPMSBORLSB = 3,
  RSP_FFTW_REG_BUSY = 4 /* RO */, RSP_FFTW_REG_OVERFLOW = 5 /* RO */ };
enum { /* AXI4LogMagMuxBlock (generators/logMagMux/src/main/scala/LogMagMuxDspBlock.scala:31-39) */
  RSP_MAG_REG_SEL = 0 /* 0 bypass, 1 |z|^2, 2 JPL magnitude, 3 log2 magnitude */ };
enum { /* AXI4CFARBlock (generators/cfar/src/main/scala/CFARDspBlock.scala:96-130) */
  RSP_CFAR_REG_FFTWIN = 0, RSP_CFAR_REG_THRESHOLDSCALER = 1, RSP_CFAR_REG_LOGORLINEARMODE = 2, RSP_CFAR_REG_DIVSUM = 3,
  RSP_CFAR_REG_PEAKGROUPING = 4, RSP_CFAR_REG_CFARALGORITHM = 5, RSP_CFAR_REG_CFARMODE = 6, RSP_CFAR_REG_WINDOWCELLS = 7,
  RSP_CFAR_REG_GUARDCELLS = 8, RSP_CFAR_REG_INDEXLAGG = 9, RSP_CFAR_REG_INDEXLEAD = 10, RSP_CFAR_REG_SUBWINDOWSIZE = 11 };
enum { /* PLFGDspBlockMem (generators/plfg/src/main/scala/PLFGDspBlockMem.scala:182-217) */
  RSP_PLFG_REG_ENABLE = 0, RSP_PLFG_REG_RESET = 1, RSP_PLFG_REG_FRAMENUM = 2, RSP_PLFG_REG_INTERFRAMENUMOFSAMPLES = 3,
  RSP_PLFG_REG_DIFFERENTCHIRPSNUM = 4, RSP_PLFG_REG_STARTINGPOINT = 5, RSP_PLFG_REG_SEGMENTNUMS0 = 6
  /* then segmentNums[maxChirpOrdinalNum], repeatedChirpNums[maxNumOfDifferentChirps], chirpOrdinalNums[maxNumOfDifferentChirps] */ };
enum { /* AXI4NCOLazyModuleBlock with pincType Streaming, poffType Fixed and useMultiplier (NCOLazyModuleBlock.scala:262-269); without
        * the multiplier this configuration has no register file */
  RSP_NCO_REG_ENABLEMULTIPLYING = 0, RSP_NCO_REG_MULTIPLYINGFACTOR = 1 /* tableWidth bits, read as FixedPoint(w, w - 2) */ };
/* fp32 chains only: the threshold scaler as a float (the FixedPoint register has no fp32 meaning) */
#define RSP_CFAR_REG_THRESHOLDSCALER_F32 0x100u

typedef struct rsp_chain rsp_chain;

/* one detection = one output beat with peak = 1, unpacked */
typedef struct {
  uint32_t chirp;      /* index of the chirp (row) inside the call: rx * n_chirps + chirp */
  uint32_t bin;        /* fftBin                                                          */
  int32_t threshold;   /* fixed: raw protoThreshold bits; fp32: float bits                 */
  int32_t cut;         /* fixed: raw protoIn bits of the cell under test; fp32: float bits  */
} rsp_detection;

/* device-side detection list the caller provides: `list` has room for `capacity` entries, `count` is one
 * uint32 the kernels atomically advance (the caller zeroes it; it may exceed capacity = overflow, entries
 * past capacity are dropped).  Both are device pointers. */
typedef struct {
  rsp_detection *list;
  uint32_t *count;
  uint32_t capacity;
} rsp_dets;

/* ---- life cycle ------------------------------------------------------------------------------------------
 * replaces: LazyModule(new RspChainVanilla(params)) / new FftMagCfarChainVanilla(params)
 *           (src/main/scala/RspChain.scala:35-61,150; src/main/scala/FftMagCfarChain.scala:31-49) */
rsp_status rsp_chain_create(const rsp_chain_params *params, int device, rsp_chain **out);
void rsp_chain_destroy(rsp_chain *c);

/* replaces: AXI4MasterModel.memWriteWord(blockBase + 4*word_index, value) / memReadWord
 *           (src/test/scala/FftMagCfarChainTester.scala:82-132; regmaps cited above) */
rsp_status rsp_reg_write(rsp_chain *c, rsp_block blk, uint32_t word_index, uint32_t value);
rsp_status rsp_reg_read(rsp_chain *c, rsp_block blk, uint32_t word_index, uint32_t *value);
/* replaces: writes into the PLFG config RAM at plfgRAM + 4*word_index, word = {reset, sign, nSamples, slope}
 *           (generators/plfg/src/main/scala/PLFGDspBlockMem.scala:108-128,168-178) */
rsp_status rsp_plfg_ram_write(rsp_chain *c, uint32_t word_index, uint32_t value);
/* WindowFunctionTypes.User(dataWidth, dv) (WindowFunc.scala:30-33): the numPoints doubles of a user-specified window; the chain
 * must have been created with windowFunc = RSP_WINDOW_USER (until this call its window is all ones).  n must equal numPoints
 * (SDFFFT.scala:59: "Length of specified window function is not the same as fft size").  Synchronises the device.            */
rsp_status rsp_fft_window_user(rsp_chain *c, const double *coeffs, uint32_t n);
/* replaces: memWriteWord(ramAddress.base + word_index * beatBytes, value) into the window RAM of FFTBlockWithWindowing
 *           (FFTBlockWithWindowing.scala:44-63; generators/sdf-fft/src/test/scala/FFTBlockWithWindowingSpec.scala:104-108): the low
 *           windowWidth bits of value are one protoWin coefficient, FixedPoint(windowWidth, windowWidth - 2).  The RAM starts as zeros.
 *           Takes effect with the next run call. */
rsp_status rsp_fft_window_ram_write(rsp_chain *c, uint32_t word_index, uint32_t value);

/* ---- the hot path, device buffers ------------------------------------------------------------------------
 * replaces: streaming n_rx*n_chirps chirps of fftSize beats into FftMagCfarChainVanilla.streamNode and reading
 *           the CFAR beats back (src/main/scala/FftMagCfarChain.scala:47; tester hot loop
 *           src/test/scala/FftMagCfarChainTester.scala:137-151).
 * d_in : fixed  -> uint32 beats Cat(re16, im16) (generators/sdf-fft/src/main/scala/FFTBlock.scala:106)
 *        fp32   -> float2 (re, im)
 * d_out: fixed  -> uint32 beats Cat(threshold, fftBin, peak) or the CUT when sendCut
 *                  (generators/cfar/src/main/scala/CFARDspBlock.scala:78-90); may be NULL
 *        fp32   -> float thresholds, one per bin; may be NULL
 * d_peak_mask: fp32 only (NULL otherwise / when not wanted): 1 bit per bin, bin n of chirp c at bit (n & 31)
 *        of word c*N/32 + n/32
 * dets : optional compacted detection list (NULL = none)
 * The transform size is the run-time fftSize register (N = 1 << fftSize). */
rsp_status rsp_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                   const rsp_dets *dets, void *stream);

/* replaces: RspChainVanilla with the PLFG enabled (src/main/scala/RspChain.scala:57-58;
 *           src/test/scala/RspChainVanillaTester.scala:86-147): the chirps are synthesised on the device by the
 *           PLFG -> NCO kernels from the registers / config RAM and consumed by the same fused chain.
 * d_nco_out (optional) receives the NCO beats Cat(cos16, sin16) (generators/nco/src/main/scala/NCOLazyModuleBlock.scala:452). */
rsp_status rsp_run_from_generator(rsp_chain *c, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                                  const rsp_dets *dets, uint32_t *d_nco_out, void *stream);
/* rsp_run_from_generator for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  No counterpart in the reference.  Same results as rsp_run_from_generator (without the
 * d_nco_out option), but the NCO beats between the generator and the chain kernel, and the PLFG segment list, live in two file-scope device
 * arrays of fixed size instead of buffers the handle allocates: the call performs no cudaMalloc / cudaFree.  Limits: n_rx * n_chirps *
 * fftSize <= RSP_STATIC_SCRATCH_BEATS samples and <= RSP_STATIC_MAX_SEGMENTS PLFG segments per period; beyond them RSP_EINVAL (the message
 * names the limit).  Every GPU that loads the library holds the arrays (4 MiB + 16 KiB) until exit.  All calls on a GPU share them: calls that
 * overlap on the GPU (other streams, other handles) race; the arrays keep the last call's contents and nothing in them needs to start at 0. */
#define RSP_STATIC_SCRATCH_BEATS (1u << 20)
#define RSP_STATIC_MAX_SEGMENTS 512
rsp_status rsp_run_from_generator_static(rsp_chain *c, uint32_t n_chirps, uint32_t n_rx, void *d_out, uint32_t *d_peak_mask,
                                         const rsp_dets *dets, void *stream);

/* ---- single blocks (each reference DspBlock is usable on its own) ----------------------------------------- */
/* replaces AXI4FFTBlock (generators/sdf-fft/src/main/scala/FFTBlock.scala:122-124): beats in -> beats out */
rsp_status rsp_fft_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, void *d_out, void *stream);
/* Wide FFT block (dataWidth > 16, twiddleWidth > 16 or expandLogic != 0; radix 2^2 DIF; numPoints <= 8192): beats are 64 bits wide
 * in (the low 2 * dataWidth bits are Cat(re, im), `in.bits.data.asTypeOf(params.protoIQ)`, FFTBlock.scala:106) and
 * rsp_fft_out_beat_bytes() wide out: Cat(re, im) of the last stage's type, dataWidth + sum(expandLogic) bits per half
 * (`fft.io.out.bits.asUInt`, :113), one uint64, or two (low word first) when that needs more than 64 bits.  rsp_run (the chain) refuses
 * such a configuration: LogMagMux and CFAR take 16-bit cells. */
uint32_t rsp_fft_in_beat_bytes(rsp_chain *c);    /* 4 or 8 */
uint32_t rsp_fft_out_beat_bytes(rsp_chain *c);   /* 4, 8 or 16 */
/* FFTParams.overflowReg = 1: the per-stage overflow status of the handle's most recent rsp_fft_run / rsp_run call on `stream`
 * (waits for it): bit i = some butterfly of PHYSICAL stage i produced a sum or difference at or above 2^(w - 1), w the width of the
 * stage's input type (`tooBig`, SDFChainRadix22.scala:473-481; the check looks at the value before div2, so it also reports stages
 * that scale).  The hardware register reloads every clock; this is its OR over the call.  RSP_FFT_REG_OVERFLOW reads the same word
 * through the reference's log2Ceil(numStages)-bit field (FFTBlock.scala:67,97), i.e. its low bits only. */
rsp_status rsp_fft_overflow_flags(rsp_chain *c, uint32_t *mask, void *stream);
/* replaces AXI4LogMagMuxBlock (generators/logMagMux/src/main/scala/LogMagMuxDspBlock.scala:59-61); n_beats beats */
rsp_status rsp_mag_run(rsp_chain *c, const void *d_in, uint64_t n_beats, void *d_out, void *stream);
/* replaces AXI4CFARBlock (generators/cfar/src/main/scala/CFARDspBlock.scala:134-136): n_chirps rows of fftWin beats
 * (fixed: low inWidth bits of each beat are the sample, :49; fp32: floats) */
rsp_status rsp_cfar_run(rsp_chain *c, const void *d_in, uint32_t n_chirps, void *d_out, uint32_t *d_peak_mask,
                        const rsp_dets *dets, void *stream);
/* replaces PLFGDspBlockMem -> AXI4NCOLazyModuleBlock (src/main/scala/RspChain.scala:57): n_samples NCO beats
 * Cat(cos16, sin16); d_plfg_out (optional) receives the PLFG words (one int32 per sample, sign-extended) */
rsp_status rsp_generate(rsp_chain *c, uint64_t n_samples, uint32_t *d_nco_out, int32_t *d_plfg_out, void *stream);

/* ---- the hot path, host buffers (the call a reference user's tester makes) ----------------------------------
 * Same as rsp_run but in/out are HOST pointers; the library pipelines host->device copies, the fused kernel and
 * device->host copies over its own streams in chunks and returns when h_out is complete.  Pinned buffers from
 * rsp_host_alloc are copied directly; pageable ones are staged through internal pinned buffers.
 * h_dets/h_n_dets (optional): detections gathered on the host, up to max_dets; *h_n_dets = number found. */
rsp_status rsp_run_host(rsp_chain *c, const void *h_in, uint32_t n_chirps, uint32_t n_rx, void *h_out, uint32_t *h_peak_mask,
                        rsp_detection *h_dets, uint32_t max_dets, uint32_t *h_n_dets);
rsp_status rsp_host_alloc(void **p, size_t bytes); /* pinned host memory */
void rsp_host_free(void *p);

/* ---- multi-GPU: gathering the detection lists ------------------------------------------------------------------
 * Frames / chirps are sharded over the GPUs with no collective on the data path (rows are independent:
 * generators/cfar/src/main/scala/CFARCoreWithMem.scala:93-104); the detection lists are the only exchange.  One process per GPU.
 * replaces: nothing in the reference (one stream, one FPGA); this is SURVEY 8b's rsp_gather_detections.
 * nccl_comm is an ncclComm_t (any communicator of the process: the helpers below create one without another NCCL binding; NCCL itself
 * is bound at run time from libnccl.so.2).  Every rank passes the same slot_capacity (<= its local->capacity).  On return (asynchronous
 * on `stream`) d_counts[r] = rank r's detection count and d_gathered[r * slot_capacity + i], i < min(d_counts[r], slot_capacity), its
 * detections; d_gathered has room for n_ranks * slot_capacity entries, d_counts for n_ranks words (both on the device). */
#define RSP_NCCL_UNIQUE_ID_BYTES 128
rsp_status rsp_gather_detections(rsp_chain *c, void *nccl_comm, const rsp_dets *local, uint32_t slot_capacity,
                                 rsp_detection *d_gathered, uint32_t *d_counts, void *stream);
rsp_status rsp_nccl_get_unique_id(uint8_t id[RSP_NCCL_UNIQUE_ID_BYTES]);                 /* ncclGetUniqueId, on one rank; ship it to the others */
rsp_status rsp_nccl_comm_init_rank(void **comm, int n_ranks, const uint8_t id[RSP_NCCL_UNIQUE_ID_BYTES], int rank, int device);
void rsp_nccl_comm_destroy(void *comm);

/* ---- slow-time (Doppler) dimension, SURVEY 8(f)2 -----------------------------------------------------------------
 * replaces: nothing in the reference -- its chains stop at the 1-D range CFAR.  The reference's own blocks run along slow time as a
 * second handle whose FFT has numPoints = chirps per frame (rsp_fft_run / rsp_cfar_run on Doppler-major rows); these two calls are the
 * data movements in between, asynchronous on `stream` like every other call.
 * rsp_corner_turn: d_out[b][c][r] = d_in[b][r][c] for b < batch, r < n_rows, c < n_cols, elements of 2, 4 or 8 bytes; out of place.
 * rsp_mag_combine_run: non-coherent RX combining.  d_in holds n_rx maps of n_beats beats each, one after the other; every beat goes
 * through the handle's LogMagMux (selector 1, 2 or 3) and d_out[i] = (sum over rx of the magnitudes of beat i) >> log2(n_rx), as one
 * LogMagMux output beat Cat(0, value) (fp32: the mean as a float).  n_rx is a power of two. */
rsp_status rsp_corner_turn(rsp_chain *c, const void *d_in, void *d_out, uint32_t n_rows, uint32_t n_cols, uint32_t batch,
                           uint32_t elem_bytes, void *stream);
rsp_status rsp_mag_combine_run(rsp_chain *c, const void *d_in, uint64_t n_beats, uint32_t n_rx, void *d_out, void *stream);

/* ---- wire formats of the FPGA builds: host-side decoders, no device work ------------------------
 * replaces: generators/jtag2mm (JtagController + JTAGToMasterAXI4, jtagToMaster.scala:32-123,525-590) as the way register writes reach
 * the chain, and the AXI4StreamWidthAdapters around the UART of src/main/scala/RxFftCfarMagTxChain.scala:27-46.
 * rsp_jtag_decode turns a GPIO capture (one byte per sample; TCK / TMS / TDI on the bits given by the masks, as
 * generators/jtag2mm/src/test/python/gpio_for_jtag.py:27-52 writes them: 0x10 / 0x20 / 0x40) into the bus transactions the JTAG master
 * issues: instructions 0x1 write, 0x2 address (upper half of the data register), 0x3 data, 0x4 read, 0x8 burst count, 0x9 burst write,
 * 0xA burst index, 0xB burst data, 0xC burst read (README "User manual").  The caller applies them (memWriteWord(base + addr, data):
 * rsp_reg_write / rsp_plfg_ram_write after its own address decode).  Returns RSP_ENOMEM with *n_ops = the number needed when ops_cap is
 * too small.  rsp_uart_to_beats / rsp_beats_to_uart regroup a byte stream into stream beats and back (beats are native little-endian
 * words in memory); msb_first selects which byte of a beat travels first -- the adapter's source is not in the reference tree, so this is
 * the caller's statement about the capture (0 = least significant byte first). */
typedef enum rsp_bus_kind { RSP_BUS_WRITE = 0, RSP_BUS_READ = 1 } rsp_bus_kind;
typedef struct rsp_bus_op {
  uint32_t kind;             /* rsp_bus_kind */
  uint32_t addr;             /* addressValue (+ beatBytes per burst element): UInt((beatBytes * 4).W), jtagToMaster.scala:534 */
  uint32_t data_lo, data_hi; /* the data word of a write (data_hi only with beatBytes = 8) */
} rsp_bus_op;
rsp_status rsp_jtag_decode(const uint8_t *samples, size_t n_samples, uint8_t tck_mask, uint8_t tms_mask, uint8_t tdi_mask,
                           uint32_t ir_length, uint32_t beat_bytes, uint32_t burst_max, rsp_bus_op *ops, uint32_t ops_cap,
                           uint32_t *n_ops);
rsp_status rsp_uart_to_beats(const uint8_t *bytes, size_t n_bytes, uint32_t beat_bytes, int msb_first, void *beats, size_t *n_beats);
rsp_status rsp_beats_to_uart(const void *beats, size_t n_beats, uint32_t beat_bytes, int msb_first, uint8_t *bytes);

/* ---- introspection ------------------------------------------------------------------------------------------ */
const char *rsp_status_str(rsp_status s);
const char *rsp_last_error(void);
int rsp_abi_version(void);
/* kernels launched by this handle since creation (bench.py reports it as gpu_launches) */
uint64_t rsp_kernel_launches(const rsp_chain *c);
/* name of the kernel variant the current registers select for rsp_run, e.g. "chain_fixed<10,fast>" */
const char *rsp_variant_name(rsp_chain *c);
/* algorithmic HBM bytes one chirp of the current configuration moves through rsp_run (in + out) */
uint64_t rsp_bytes_per_chirp(rsp_chain *c);

#ifdef __cplusplus
}
#endif
#endif
