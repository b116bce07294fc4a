/*
 * rsp_oracle.h -- CPU oracle for the rsp-chains hot path (TEST INFRASTRUCTURE ONLY).
 *
 * This is a plain-C restatement of the arithmetic of the reference's Chisel generators
 * (PLFG -> NCO -> SDF-FFT -> LogMagMux -> CFAR).  It is the checker the CUDA path is held to.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it.  The product library (rsp_chains_b200/csrc) never links, loads or calls it.
 *
 * PARITY STATUS: "parity unpinned" at the bit level.  The reference (Scala/Chisel) cannot be
 * elaborated or simulated here (no JVM) and ships no golden vectors; its arithmetic bottoms out in
 * an un-vendored fork of dsptools (edu.berkeley.cs:rocket-dsptools:1.2-SNAPSHOT, build.sbt:42) whose
 * semantics are restated from its published source as recalled ("Assumed semantics").
 * What IS pinned (tests/test_oracle_*.py): PLFG known-answer vectors (PLFGSpec.scala:42-62) and the
 * reference's own Python PLFG() outputs (exact); every tolerance criterion the reference's own tests
 * apply to FFT / JPL / CFAR / NCO; the published SQNR curves; the Scala golden CFAR models restated
 * line by line in fp64.
 *
 * All file:line citations are relative to /root/reference/.
 */
#ifndef RSP_ORACLE_H
#define RSP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* dsptools TrimType (the reference fork has "eleven trim types", generators/sdf-fft/README.md:100;
 * names from generators/sdf-fft/src/test/scala/Radix22SQNRSpec.scala:184). */
enum {
  RSPO_TRIM_ROUND_HALF_UP = 0,          /* ties toward +inf            */
  RSPO_TRIM_FLOOR = 1,                  /* = RoundDown                 */
  RSPO_TRIM_CONVERGENT = 2,             /* = RoundHalfToEven           */
  RSPO_TRIM_ROUND_UP = 3,               /* ceil                        */
  RSPO_TRIM_ROUND_TOWARDS_ZERO = 4,
  RSPO_TRIM_ROUND_TOWARDS_INFINITY = 5, /* away from zero              */
  RSPO_TRIM_ROUND_HALF_DOWN = 6,        /* ties toward -inf            */
  RSPO_TRIM_ROUND_HALF_TOWARDS_ZERO = 7,
  RSPO_TRIM_ROUND_HALF_TOWARDS_INFINITY = 8,
  RSPO_TRIM_ROUND_HALF_TO_ODD = 9,
  RSPO_TRIM_NO_TRIM = 10
};

/* FFTParams.fixed subset that changes values (generators/sdf-fft/src/main/scala/SDFFFTParams.scala:90-149) */
typedef struct {
  int32_t log_n;          /* log2(numPoints) actually transformed (run-time fftSize register) */
  int32_t data_width;     /* dataWidth (16 in every chain config)                             */
  int32_t bin_point;      /* binPoint                                                         */
  int32_t twiddle_width;  /* twiddleWidth -> Q2.(tw-2)                                        */
  int32_t trim_type;      /* trimType used by div2 / context_* / trimBinary                   */
  int32_t use4muls;       /* complexUse4Muls                                                  */
  int32_t bit_reverse;    /* useBitReverse: 1 -> natural-order output                         */
  uint32_t keep_msb_mask; /* bit i = keepMSBorLSB(i) for stage i of THIS transform (1 = scale, 0 = keep LSB / wrap) */
  int32_t fft_dir;        /* 1 = FFT, 0 = IFFT (SDFChainRadix22.scala:205-211,362-383)        */
  int32_t sdf_radix;      /* 2 = sdfRadix "2" (SDFChainRadix2.scala); 0 or 4 = "2^2"          */
  int32_t decim_dit;      /* 1 = DITDecimType, 0 = DIFDecimType                               */
  int32_t num_stages;     /* log2(numPoints) of the ELABORATED chain (radix-2^2 DIT: which stages carry twiddles depends on its
                             parity, SDFChainRadix22.scala:229); 0 = same as log_n                                        */
  uint32_t expand_mask;   /* bit i = expandLogic(i) of stage i of THIS transform: the stage keeps the grown butterfly outputs in a type one
                             bit wider than its input (SDFFFTParams.scala:119-123, SDFChainRadix22.scala:453-457); radix 2^2 DIF only   */
  int32_t base_growth;    /* bits the stages in front of THIS transform's first stage have already added to the types (run-time size) */
} rspo_fft_cfg;
/* width of the output type of stage i of this transform: dataWidth + base_growth + sum expandLogic(0..i) */
int rspo_fft_stage_width(const rspo_fft_cfg *cfg, int stage);

/* MAGParams (generators/logMagMux/src/main/scala/LogMagMux.scala:14-49) + run-time sel */
typedef struct {
  int32_t data_width, bin_point;        /* proto                                 */
  int32_t data_width_log, bin_point_log;/* protoLog                              */
  int32_t log2_lookup_width;
  int32_t sel;                          /* 0 bypass, 1 |z|^2, 2 JPL, 3 log2      */
  int32_t ctx_trim;                     /* DspContext default trimType */
} rspo_mag_cfg;

/* CFARParams + run-time registers (generators/cfar/src/main/scala/CFARParams.scala:24-49,
 * CFARDspBlock.scala:96-130) */
enum { RSPO_CFAR_CORE_MEM = 0, RSPO_CFAR_CORE_LIS = 1, RSPO_CFAR_CORE_ASR = 2,
       RSPO_CFAR_CORE_OS = 3 /* integer form of the golden cfarOS (CFARUtils.scala:159-196): combined window, ascending index floor(1.5 W) */ };
typedef struct {
  int32_t core;              /* which CFARCore variant is elaborated (CFARcore.scala:62-64)      */
  int32_t in_width, in_bp;   /* protoIn                                                          */
  int32_t thr_width, thr_bp; /* protoThreshold                                                   */
  int32_t scaler_width, scaler_bp; /* protoScaler                                                */
  int32_t lead_lagg_window_size;   /* compile-time max window (sets sum register width)          */
  int32_t guard_window_size;       /* compile-time max guard                                     */
  int32_t fft_size;          /* compile-time max fft size (width of the fftBin field)            */
  int32_t min_sub_window_size;
  int32_t send_cut;
  /* run-time registers */
  int32_t fft_win;
  int32_t threshold_scaler;  /* raw bits of protoScaler */
  int32_t log_or_linear;     /* 1 = linear (multiply), 0 = log (add) */
  int32_t div_sum;
  int32_t peak_grouping;
  int32_t cfar_algorithm;    /* Lis core: 1 = GOS, 0 = CA sums */
  int32_t cfar_mode;         /* 0 CA, 1 GO, 2 SO, 3 CASH(ASR) */
  int32_t window_cells, guard_cells, index_lagg, index_lead, sub_cells;
  int32_t ctx_trim;          /* DspContext default trimType for the threshold multiply */
  int32_t edge_mode;         /* 0 = as the elaborated core does; (reserved) */
} rspo_cfar_cfg;

/* ---- primitives ---- */
int64_t rspo_trim(int64_t v, int shift, int mode);
int64_t rspo_wrap(int64_t v, int width);

/* ---- tables ---- */
/* twiddles of one multiplier stage: entries t in [0, 3*stage_n/4), (cos, -sin)(2 pi t / stage_n) in Q2.(tw-2)
 * (SDFChainRadix22.scala:255-267) */
void rspo_twiddle_table(int stage_n, int tw_width, int32_t *re, int32_t *im);
/* radix-2 stage ROM of `delay` entries (SDFChainRadix2.scala:300-306) */
void rspo_twiddle_table_radix2(int delay, int tw_width, int32_t *re, int32_t *im);
/* sine quarter table (NCOTable.scala:254-258) */
void rspo_nco_table(int table_size, int table_width, int32_t *tab);
/* log2 LUT (LogMagMux.scala:120-131) */
void rspo_log2_table(int lookup_width, int bp_log, int32_t *tab);

/* ---- FFT ---- */
/* radix-2^2 DIF, scale-or-wrap per stage; in/out raw fixed-point integers (SDFChainRadix22.scala:224-352,410-505) */
/* as rspo_fft_fixed; *overflow_mask (may be NULL) receives bit i = some butterfly output of stage i of this transform had the
 * grown sum / difference at or above 2^(w-1), w = the stage's input width: `tooBig` of SDFChainRadix22.scala:473-481 (`tooSmall`
 * can never fire: an arithmetic right shift of a negative value is never 0); grow stages report 0 (:455-457) */
int rspo_fft_fixed_ex(const rspo_fft_cfg *cfg, const int32_t *in_re, const int32_t *in_im, int32_t *out_re, int32_t *out_im,
                      uint32_t *overflow_mask);
int rspo_fft_fixed_words64(const rspo_fft_cfg *cfg, const uint64_t *in, uint64_t *out, int n_chirps, uint32_t *overflow_mask);
int rspo_fft_fixed_words_ovf(const rspo_fft_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps, uint32_t *overflow_mask);
int rspo_fft_fixed(const rspo_fft_cfg *cfg, const int32_t *in_re, const int32_t *in_im,
                   int32_t *out_re, int32_t *out_im);
/* the DspReal "golden" flavour of the same dataflow: every stage halves, twiddles exact */
int rspo_fft_f64(int log_n, int bit_reverse, const double *in_re, const double *in_im, double *out_re, double *out_im);
/* packed 32-bit beats Cat(re16, im16) (FFTBlock.scala:106,114), n_chirps transforms */
int rspo_fft_fixed_words(const rspo_fft_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps);

/* ---- LogMagMux ---- */
int32_t rspo_mag_one(const rspo_mag_cfg *cfg, const int32_t *log_tab, int32_t re, int32_t im, int32_t *out_re);
int rspo_mag_words(const rspo_mag_cfg *cfg, const uint32_t *in, uint32_t *out, int64_t n);
double rspo_jpl_f64(double re, double im);

/* ---- CFAR (hardware-exact integer cores) ---- */
/* x: n = fft_win raw protoIn values; thr_full: threshold before re-shaping (for compare), may be NULL;
 * thr_out: threshold re-shaped to protoThreshold; peak: 0/1 */
int rspo_cfar_fixed(const rspo_cfar_cfg *cfg, const int32_t *x, int64_t *thr_full, int32_t *thr_out, uint8_t *peak);
/* stream words: low 16 bits of in (CFARDspBlock.scala:49) -> Cat(threshold, fftBin, peak) (CFARDspBlock.scala:90) */
int rspo_cfar_words(const rspo_cfar_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps);

/* ---- CFAR Scala golden models restated in fp64 (CFARUtils.scala:29-272); thr arrays sized by caller,
 * returned lengths follow the golden's own (quirky) edge padding; peaks may contain duplicates exactly as the golden */
int rspo_golden_cfar_ca(const double *sig, int n, int mode /*0 CA,1 GO,2 SO*/, int ref_cells, int guard_cells,
                        int consider_edges, double scaling, int log_mode, double *thr, int *n_thr, int *peaks, int *n_peaks);
int rspo_golden_cfar_os(const double *sig, int n, int ref_cells, int guard_cells, double scaling, int log_mode,
                        double *thr, int *n_thr, int *peaks, int *n_peaks);
int rspo_golden_cfar_gos(const double *sig, int n, int ref_cells, int guard_cells, int index_lagg, int index_lead,
                         int mode, int consider_edges, double scaling, int log_mode,
                         double *thr, int *n_thr, int *peaks, int *n_peaks);
int rspo_golden_cfar_cash(const double *sig, int n, int ref_cells, int sub_cells, int guard_cells, double scaling,
                          int log_mode, double *thr, int *n_thr, int *peaks, int *n_peaks);

/* fp64 chain flavour ("DspReal"): thresholds by the hardware edge policy but in doubles */
int rspo_cfar_f64(const rspo_cfar_cfg *cfg, double scaler, const double *x, double *thr, uint8_t *peak);

/* ---- PLFG / NCO ---- */
typedef struct {
  int32_t max_num_of_segments;     /* PLFGParams.maxNumOfSegments (RAM row stride, PLFGDspBlockMem.scala:383) */
  int32_t max_num_of_samples_width;
  int32_t output_width;            /* outputWidthInt + outputWidthFrac */
  int32_t frame_num;               /* 0 = forever (bounded by n_out)   */
  int32_t interframe_num_of_samples;
  int32_t different_chirps_num;
  int32_t starting_point;
  const uint32_t *segment_nums;        /* [maxChirpOrdinalNum]        */
  const uint32_t *repeated_chirp_nums; /* [maxNumOfDifferentChirps]   */
  const uint32_t *chirp_ordinal_nums;  /* [maxNumOfDifferentChirps]   */
  const uint32_t *ram;                 /* config RAM words {reset,sign,nSamples,slope} (PLFGDspBlockMem.scala:168-178) */
  int32_t output_width_int;        /* outputWidthInt: the RAM word is sliced at outputWidthInt (point4, :161-171), the slope field and
                                      startingPoint are zero-extended into the (Int + Frac)-bit protoOut by asTypeOf; 0 = output_width */
  int32_t pause_model;             /* inter-frame pause.  0 = the RTL counter tree (:242-275): framePauseCounter starts on the tick of the
                                      frame's last sample and the frame restarts when fpc + 1 >= interframeNumOfSamples -> inter - 1
                                      pause samples (none for inter <= 1).  1 = plfg_script.py:50-53 (inter pause samples).  The two
                                      models of the reference disagree; its value-checked tests use frameNum = 1 only. */
} rspo_plfg_cfg;
/* returns number of samples produced (<= n_out); out = raw output_width-bit values, sign-extended */
int64_t rspo_plfg(const rspo_plfg_cfg *cfg, int32_t *out, int64_t n_out);

typedef struct {
  int32_t table_size, table_width, phase_width;
  int32_t phase_acc_enable;
  uint32_t phase0;     /* accumulator state before the first sample */
  uint32_t poff;
  /* output multiplier (FixedNCOParams.useMultiplier, NCOLazyModuleBlock.scala:354-364): sin and cos are each multiplied by
   * factor = enableMultiplying ? multiplyingFactor : 1 << (beatBytes * 4 - 2), both read as FixedPoint(w, w - 2), context_* with
   * binaryPointGrowth 0 under roundingMode, result reinterpreted as protoOut (low w bits) */
  int32_t use_multiplier, enable_multiplying;
  uint32_t multiplying_factor;
  int32_t rounding_mode;   /* rspo trim type */
  int32_t rasterized;      /* NCOTableRasterizedMode (NCOTable.scala:322-464): table_size need not be a power of two, phase_width = log2Ceil(table_size) + 2 */
} rspo_nco_cfg;
/* freq: PLFG words; out: Cat(cos16, sin16) beats (NCOLazyModuleBlock.scala:452); returns final phase */
uint32_t rspo_nco_words(const rspo_nco_cfg *cfg, const int32_t *freq, uint32_t *out, int64_t n);

/* ---- whole chain, fixed point: words in -> CFAR words out; n_threads > 1 uses OpenMP over chirps ---- */
typedef struct {
  rspo_fft_cfg fft;
  rspo_mag_cfg mag;
  rspo_cfar_cfg cfar;
} rspo_chain_cfg;
int rspo_chain_fixed_words(const rspo_chain_cfg *cfg, const uint32_t *in, uint32_t *out, int n_chirps, int n_threads,
                           uint32_t *fft_out /*nullable*/, uint32_t *mag_out /*nullable*/);
int rspo_max_threads(void);

/* java.util.Random (= scala.util.Random) so the reference's seeded test signals can be regenerated
 * (RspChainTesterUtils.scala:56-91, CFARCATester.scala:231-244) */
typedef struct { uint64_t seed; } rspo_jrand;
void rspo_jrand_seed(rspo_jrand *r, int64_t seed);
double rspo_jrand_next_double(rspo_jrand *r);

#ifdef __cplusplus
}
#endif
#endif
