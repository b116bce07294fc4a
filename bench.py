#!/usr/bin/env python
"""bench.py -- the headline benchmark: end-to-end chain Msamples/s (FFT -> LogMagMux -> CFAR), fixed point.

A step = one pass of the fused chain over one frame of BASELINE.json configs[4] per GPU ("streaming sweep": 8 RX x 4096 chirps
x 16384-point fixed-point range FFT + JPL magnitude + CA-CFAR): 2^29 samples, 2 GiB of beats in + 2 GiB out per GPU per step.
With N GPUs the job is N frames, sharded one frame per rank (weak scaling, no data-path collective); NCCL only gathers the
detection lists, and that gather is timed on its own (`gather_ms`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  torchrun ... bench.py --gpus N ...      one rank per GPU

Prints ONE JSON line (rank 0).  `value` is device-resident throughput, `e2e` goes through rsp_run_host with pinned host buffers
(H2D and D2H inside the timed region; `e2e.detections_only` is the same call with h_out = NULL, detections as the only result),
`roofline` relates the fused kernel to the measured HBM peak, `configs` holds one line per BASELINE configuration (C1..C4, the fp32
and OS-CFAR kernels, the chain fed by the PLFG/NCO generator; N = 1 only), `cpu_baseline` times the CPU oracle (a port of the
reference's arithmetic; the Scala/Chisel reference itself cannot run here) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LOG_N = 14
N_POINTS = 1 << LOG_N
N_RX = 8
CHIRPS_PER_FRAME = 4096
FRAMES_PER_GPU = 1                # frames per step per GPU: 32768 rows = 2^29 samples, 2 GiB in + 2 GiB out
DISTINCT_ROWS = 512               # synthesised on the host per frame, tiled to the frame on the device
BYTES_PER_SAMPLE = 8              # Cat(re16, im16) in + Cat(threshold, bin, peak) out  
REGS = dict(fftWin=N_POINTS, thresholdScaler=4, logOrLinearMode=1, divSum=6, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=8)
WORKLOAD = (f"{FRAMES_PER_GPU} frame(s) x {N_RX} RX x {CHIRPS_PER_FRAME} chirps x {N_POINTS}-pt fixed-point (Q16.0) range FFT + JPL mag + "
            f"CA-CFAR (guard 8, ref 64) per GPU per step [BASELINE configs[4]]")
METRIC = "end-to-end chain Msamples/s"


def fixed_params(n, W, G, ctype=None):
    import rsp_chains_b200 as R
    from rsp_chains_b200.params import CFARParams, FFTParams, FftMagCfarVanillaParameters, FixedPoint, MAGParams
    return FftMagCfarVanillaParameters(
        fftParams=FFTParams.fixed(dataWidth=16, twiddleWidth=16, numPoints=n, useBitReverse=True, runTime=True, binPoint=0),
        magParams=MAGParams.fixed(dataWidth=16, binPoint=0, dataWidthLog=16, binPointLog=9, log2LookUpWidth=9),
        cfarParams=CFARParams(protoIn=FixedPoint(16, 0), protoThreshold=FixedPoint(16, 0), protoScaler=FixedPoint(16, 0),
                              leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=ctype or R.CACFARType, sendCut=False))


def f32_params(n, W, G):
    import rsp_chains_b200 as R
    from rsp_chains_b200.params import CFARParams, FFTParams, FftMagCfarVanillaParameters, MAGParams
    return FftMagCfarVanillaParameters(fftParams=FFTParams.DSPReal(numPoints=n, useBitReverse=True, runTime=True), magParams=MAGParams.fixed(),
                                       cfarParams=CFARParams(leadLaggWindowSize=W, guardWindowSize=G, fftSize=n, CFARAlgorithm=R.CACFARType, sendCut=False))


def chain_params():
    return fixed_params(N_POINTS, 64, 8)


def ncu_traffic():
    """DRAM bytes per launch of the headline kernel from the committed `ncu --set full` capture (profiles/), or None.
    The capture is of the same command (one C5 frame per launch), so it is per launch like `achieved`."""
    import glob
    best = None
    pats = ("r01_i_ncu_full_fx14_kernel.txt", "r02*_ncu_full_fx14*.txt")
    files = [f for p in pats for f in sorted(glob.glob(os.path.join(ROOT, "profiles", p)))]
    for f in files:
        rd = wr = None
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for line in open(f):
            c = line.rstrip("\n").split("\t")
            if len(c) == 3 and c[0] == "dram__bytes_read.sum" and c[1] in unit:
                rd = float(c[2]) * unit[c[1]]
            if len(c) == 3 and c[0] == "dram__bytes_write.sum" and c[1] in unit:
                wr = float(c[2]) * unit[c[1]]
        if rd is not None and wr is not None:
            # r01 captured 1/8 frame (4096 rows) per launch: scale to the bench launch; r02 captures are of the bench launch itself
            scale = 8.0 if os.path.basename(f).startswith("r01") else 1.0
            best = ((rd + wr) * scale, os.path.basename(f) + (" (x8: captured on 4096 rows)" if scale != 1.0 else ""))
    return best


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy burst)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.mark_at = index, [], None, 0

    def wait_first(self, timeout=8.0):
        """nvidia-smi needs 0.1 - 0.5 s to come up (longer on an 8-GPU box): wait for its first line so that the timed region is sampled from its start"""
        t0 = time.perf_counter()
        while self.proc is not None and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.01)

    def mark(self):
        """the timed region starts here: earlier rows (start-up, warm-up) are not counted"""
        self.mark_at = len(self.rows)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows[self.mark_at:]:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    """all host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which is not what the CPU arm wants)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def oracle_setup():
    from oracle import oracle as O
    from oracle.glue import oracle_cfgs
    O.build()
    fft, mag, cfar = oracle_cfgs(O, chain_params(), LOG_N, 2, REGS)
    return O, fft, mag, cfar


def time_oracle(beats, target_s=12.0):
    """all host threads, bounded sample sized to ~target_s of CPU work"""
    O, fft, mag, cfar = oracle_setup()
    threads = host_threads()
    probe = beats[:max(threads, 16)]
    t0 = time.perf_counter(); O.chain_fixed_words(fft, mag, cfar, probe, n_threads=threads); dt = time.perf_counter() - t0
    rate = probe.size / max(dt, 1e-9)
    rows = int(max(probe.shape[0], (rate * target_s) // N_POINTS))
    reps = max(1, -(-rows // beats.shape[0]))
    sample = beats[:min(rows, beats.shape[0])]
    t0 = time.perf_counter()
    for _ in range(reps):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    dt = time.perf_counter() - t0
    return (sample.size * reps / dt / 1e6, threads,
            f"{reps} x {sample.shape[0]} chirps x {N_POINTS} of the same synthetic workload, {dt:.1f} s wall, OpenMP over chirps")


def host_rows(frame: int):
    """DISTINCT_ROWS chirps of frame `frame` (its own seed: every rank / frame processes different data)"""
    from rsp_chains_b200 import synth
    # noise 2^-2 of full scale: after the 1/N of the scale-every-stage transform the noise floor of a 16384-point row is ~16 LSB
    # (at the 2^-7 of the 1024-point recipe it would quantise to 0 and every second bin would be a "detection")
    return synth.chirps_fixed(DISTINCT_ROWS, N_POINTS, seed=11110 + frame, noise=2.0 ** -2)


def run_reference(args):
    """the reference arm: the CPU implementation of the path (oracle port; the Scala/Chisel reference cannot run here)"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O, fft, mag, cfar = oracle_setup()
    threads = host_threads()
    base = host_rows(0)
    # each step: a bounded sample of the workload (~1.5 s of CPU work)
    probe = base[:max(threads, 16)]
    t0 = time.perf_counter(); O.chain_fixed_words(fft, mag, cfar, probe, n_threads=threads); dt = time.perf_counter() - t0
    rows = int(min(base.shape[0], max(probe.shape[0], (probe.shape[0] * 1.5 / max(dt, 1e-9)))))
    sample = base[:rows]
    for _ in range(args.warmup):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    dt = time.perf_counter() - t0
    val = sample.size * args.steps / dt / 1e6
    emit({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int16 fixed point (Q16.0, int64 intermediates)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{rows} chirps x {N_POINTS} per step on the host"},
        "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": threads, "kind": "port",
                         "sample": f"{rows} chirps x {N_POINTS} per step, {args.steps} steps"},
        "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is a Scala/Chisel hardware generator (no JVM here): CPU arm = oracle/ C port of its arithmetic, OpenMP over chirps",
    })


# ---------------------------------------------------------------------------------------------------------------------
# one line per BASELINE configuration (N = 1): device-resident, inputs + outputs larger than L2 ("batched"), plus the time of
# the configuration as BASELINE states it (one frame; small frames are L2-resident and launch-latency bound -- flagged)
def _timed(torch, fn, steps=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]))


def config_lines(torch, peak):
    import rsp_chains_b200 as R
    from rsp_chains_b200 import synth
    out = []
    ca = lambda n, W, G, ds, mode=0: dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=ds, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)  # noqa: E731

    def case(name, params, log_n, regs, frame_rows, n_rx, frames, fp32=False, bps=8.0, note=None, dense=False):
        n = 1 << log_n
        rows = frame_rows * frames
        distinct = min(64, rows)
        if fp32:
            base = torch.from_numpy(synth.chirps_f32(distinct, n, seed=1)).cuda()
            d_in = base.repeat((rows + distinct - 1) // distinct, 1, 1)[:rows].contiguous()
            d_out = torch.empty((rows, n), dtype=torch.float32, device="cuda")
            d_mask = torch.zeros((rows, n // 32), dtype=torch.int32, device="cuda")
        else:
            beats = synth.random_beats(distinct, n, seed=1) if dense else synth.chirps_fixed(distinct, n, seed=1)
            base = torch.from_numpy(beats.view(np.int32)).cuda()
            d_in = base.repeat((rows + distinct - 1) // distinct, 1)[:rows].contiguous()
            d_out = torch.empty_like(d_in)
            d_mask = None
        with R.FftMagCfarChainVanilla(params) as ch:
            ch.configure(log_n, 2, regs)
            if fp32:
                ch.set_threshold_scaler_f32(3.5)
            ms = _timed(torch, lambda: ch.run(d_in, rows // n_rx, n_rx, out=d_out, peak_mask=d_mask))
            ms1 = ms if frames == 1 else _timed(torch, lambda: ch.run(d_in, frame_rows // n_rx, n_rx, out=d_out, peak_mask=d_mask))
            variant = ch.variant
        samples = rows * n
        gbs = samples * bps / (ms * 1e-3) / 1e9
        line = {"config": name, "variant": variant, "frames_per_launch": frames, "rows": rows, "n": n, "ms": round(ms, 4),
                "Msamples_per_s": round(samples / ms / 1e3, 1), "GBps_algorithmic": round(gbs, 1), "frac": round(gbs / peak, 4),
                "bytes_per_sample": bps, "footprint_MiB": round(samples * bps / 2 ** 20, 1), "one_frame_ms": round(ms1, 4),
                "one_frame_Msamples_per_s": round(frame_rows * n / ms1 / 1e3, 1)}
        if note:
            line["note"] = note
        out.append(line)
        del d_in, d_out
        torch.cuda.empty_cache()

    case("C1 single 1024-pt chirp, fixed CA (W32 G4)", fixed_params(1024, 64, 4), 10, ca(1024, 32, 4, 5), 1, 1, 65536,
         note="one_frame = the single chirp of configs[0] (launch-latency bound, L2-resident)")
    case("C2 one frame 256 x 1024 fixed CA (W16 G4)", fixed_params(1024, 64, 4), 10, ca(1024, 16, 4, 4), 256, 1, 256,
         note="one frame is 2 MiB (L2-resident); batched line = 256 frames per launch")
    app = R.fft_mag_cfar_chain_vanilla_app_params()
    case("FftMagCfarChainVanillaApp parameters (binPoint 12, leadLagg 64, guard 4) 256 x 1024 fixed CA", app, 10,
         dict(fftWin=1024, thresholdScaler=int(3.5 * 4096), logOrLinearMode=1, divSum=6, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4), 256, 1, 256,
         note="src/main/scala/FftMagCfarChain.scala:76-120: dsptools' 3-multiplier product with per-product trim, thresholds with binary point 12")
    for mode, nm in ((1, "GOCA"), (2, "SOCA")):
        case(f"C3 4 RX x 512 x 2048 fp32 {nm} (W32 G4)", f32_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, mode), 2048, 4, 16, fp32=True, bps=12.125,
             note="fp32 flavour; one frame is 48 MiB + 24 MiB out (L2-resident); batched = 16 frames")
    os_regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
    case("C4 OS-CFAR 1024 x 4096 fixed, reading (A): GOS core, 64 cells per side, k=48", fixed_params(4096, 64, 4, R.GOSCFARType), 12, os_regs, 1024, 1, 16,
         note="CFARCoreWithLis (indexLagg = indexLead = 48); radix selection on bit planes; one frame is 32 MiB (L2-resident)")
    case("C4 reading (A) on full-range random beats (every bit plane populated everywhere: worst case of the selection)", fixed_params(4096, 64, 4, R.GOSCFARType), 12, os_regs, 1024, 1, 16,
         dense=True, note="the bit-plane selection is data dependent; this is its slow end")
    os_b = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    case("C4 OS-CFAR 1024 x 4096 fixed, reading (B): golden cfarOS, 32 + 32 cells in one sorted window, ascending index 48", fixed_params(4096, 64, 4, R.OSGoldenCFARType), 12, os_b, 1024, 1, 16,
         note="CFARUtils.scala:159-196 (no hardware core in the reference); radix selection on bit planes (chain_fixed_fast_os)")
    out.extend(generator_lines(torch, peak))
    out.extend(cfar_only_lines(torch, peak))
    out.extend(doppler_lines(torch, peak))
    return out


def cfar_only_lines(torch, peak, n=4096, rows=16384):
    """BASELINE configs[3] as SURVEY 8d also states it: the CFAR block alone (rsp_cfar_run) on stored 16-bit magnitude profiles (exponential clutter, mean 256,
    + 4 targets), both readings of "ref window 64, k = 48"; 4 B in (a beat per cell, CFARDspBlock.scala:49) + 4 B out per cell"""
    import rsp_chains_b200 as R
    rng = np.random.default_rng(1)
    prof = np.floor(rng.exponential(256.0, (64, n))).astype(np.int32)
    prof[:, [300, 1200, 2500, 3900]] += 4096
    d_in = torch.from_numpy(prof).cuda().repeat(rows // 64, 1).contiguous()
    d_out = torch.empty_like(d_in)
    lines = []
    for name, ctype, W in (("reading (A): GOS core, 64 cells per side, k=48", R.GOSCFARType, 64), ("reading (B): golden cfarOS, 32 + 32 cells, ascending index 48", R.OSGoldenCFARType, 32)):
        regs = dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=4, indexLagg=48, indexLead=48)
        with R.FftMagCfarChainVanilla(fixed_params(n, 64, 4, ctype)) as ch:
            ch.configure(12, 2, regs)
            ms = _timed(torch, lambda: ch.cfar(d_in, rows, out=d_out))
        g = rows * n * 8.0 / (ms * 1e-3) / 1e9
        lines.append({"config": f"C4 CFAR block alone on stored magnitude profiles, 16 x 1024 x {n}, {name}", "variant": "bit-plane kernel, cfar_only", "rows": rows, "n": n,
                      "ms": round(ms, 4), "Msamples_per_s": round(rows * n / ms / 1e3, 1), "GBps_algorithmic": round(g, 1), "frac": round(g / peak, 4), "bytes_per_sample": 8.0})
    return lines


def doppler_lines(torch, peak, n_rx=8, m=1024, n=4096):
    """SURVEY 8(f)2 (not in the reference): one frame of n_rx x m chirps x n samples through RangeDopplerChain, and its corner turn alone"""
    import rsp_chains_b200 as R
    from rsp_chains_b200 import synth
    W, G = 16, 2
    regs = dict(fftWin=m, thresholdScaler=5, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=W, guardCells=G)
    base = torch.from_numpy(synth.chirps_fixed(64, n, seed=3).view(np.int32)).cuda()
    d_in = base.repeat(n_rx * m // 64, 1).view(n_rx, m, n).contiguous()
    d_out = torch.empty((n, m), dtype=torch.int32, device="cuda")
    lines = []
    with R.RangeDopplerChain(fixed_params(n, W, G), fixed_params(m, W, G), n_rx) as rd:
        rd.range.configure(n.bit_length() - 1, 2, dict(regs, fftWin=n))
        rd.doppler.configure(m.bit_length() - 1, 2, regs)
        l0 = rd.kernel_launches
        ms = _timed(torch, lambda: rd.run(d_in, out=d_out))
        launches = (rd.kernel_launches - l0)
        a, b = torch.empty_like(d_in), torch.empty_like(d_in)
        ms_t = _timed(torch, lambda: rd.range.corner_turn(d_in, m, n, a, batch=n_rx))
        ms_f = _timed(torch, lambda: rd.range.fft(d_in, n_rx * m, b))
    cells = n_rx * m * n
    bps = 8.0 + 8.0 + 8.0 + (4.0 + 4.0 / n_rx) + 8.0 / n_rx          # range FFT, corner turn, Doppler FFT, LogMagMux + RX mean, CFAR
    gbs = cells * bps / (ms * 1e-3) / 1e9
    lines.append({"config": f"range-Doppler frame {n_rx} RX x {m} chirps x {n} samples, fixed, CFAR along Doppler", "variant": "RangeDopplerChain (5 launches)",
                  "rows": n_rx * m, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(cells / ms / 1e3, 1), "GBps_algorithmic": round(gbs, 1),
                  "frac": round(gbs / peak, 4), "bytes_per_sample": round(bps, 3), "footprint_MiB": round(cells * 4 / 2 ** 20, 1),
                  "note": "SURVEY 8(f)2, not in the reference: the reference's FFT / LogMagMux / CFAR blocks run a second time along slow time; "
                          "every stage crosses HBM (no fusion across the corner turn)"})
    for name, t, b_ in (("corner turn alone (rsp_corner_turn, 4-byte beats)", ms_t, 8.0), ("range FFT block alone (rsp_fft_run)", ms_f, 8.0)):
        g = cells * b_ / (t * 1e-3) / 1e9
        lines.append({"config": f"{name}, {n_rx} x {m} x {n}", "variant": name.split(" (")[0], "rows": n_rx * m, "n": n, "ms": round(t, 4),
                      "Msamples_per_s": round(cells / t / 1e3, 1), "GBps_algorithmic": round(g, 1), "frac": round(g / peak, 4), "bytes_per_sample": b_})
    return lines


def generator_lines(torch, peak, rows=65536):
    """RspChainVanilla with the PLFG enabled (rsp_run_from_generator) and the PLFG -> NCO generator alone (rsp_generate): a triangle
    chirp of three segments, RspChainVanillaApp parameters"""
    import rsp_chains_b200 as R
    p = R.rsp_chain_vanilla_app_params()
    n, bb = 1024, p.beatBytes
    regs = dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    segs = [(1, 200, True, False), (2, 180, False, False), (0, 100, False, True)]   # (slope, nSamples, up, reset)
    W, OW = p.plfgParams.maxNumOfSamplesWidth, p.plfgParams.outputWidthInt

    def ram_word(slope, ns, up, rst):   # {reset, sign, nSamples, slope} (PLFGDspBlockMem.scala:168-178)
        return (int(rst) << (OW + W + 1)) | (int(up) << (OW + W)) | ((ns & ((1 << W) - 1)) << OW) | (slope & ((1 << OW) - 1))

    with R.RspChainVanilla(p) as ch:
        ch.configure(10, 2, regs)
        for i, sg in enumerate(segs):
            ch.memWriteWord(p.plfgRAM.base + i * bb, ram_word(*sg))
        seg0 = 6
        rep0 = seg0 + p.plfgParams.maxChirpOrdinalNum
        ord0 = rep0 + p.plfgParams.maxNumOfDifferentChirps
        for idx, v in ((2, 0), (3, 7), (4, 1), (5, 3), (seg0, len(segs)), (rep0, 2), (ord0, 0), (1, 0), (0, 1)):   # as RspChainVanillaTester.scala:86-130
            ch.memWriteWord(p.plfgAddress.base + idx * bb, v)
        d_out = torch.empty((rows, n), dtype=torch.int32, device="cuda")
        d_nco = torch.empty((rows, n), dtype=torch.int32, device="cuda")
        l0 = ch.kernel_launches
        ch.run_from_generator(rows, 1, out=d_out)
        launches_per_call = ch.kernel_launches - l0
        ms_chain = _timed(torch, lambda: ch.run_from_generator(rows, 1, out=d_out))
        ms_gen = _timed(torch, lambda: ch.generate(rows * n, d_nco))
        variant = ch.variant
    samples = rows * n
    res = []
    chain_bps = 4.0 if launches_per_call == 1 else 12.0
    for name, ms, bps, var in ((f"C1 chain fed by the PLFG -> NCO generator (RspChainVanilla), {launches_per_call} launch(es) per call", ms_chain, chain_bps, "generator + " + variant),
                               ("PLFG -> NCO generator alone", ms_gen, 4.0, "generator_table_kernel")):
        gbs = samples * bps / (ms * 1e-3) / 1e9
        res.append({"config": name, "variant": var, "rows": rows, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(samples / ms / 1e3, 1),
                    "GBps_algorithmic": round(gbs, 1), "frac": round(gbs / peak, 4), "bytes_per_sample": bps,
                    "footprint_MiB": round(samples * bps / 2 ** 20, 1)})
    return res


_JSON_FD = None


def claim_stdout():
    """stdout carries the ONE JSON line and nothing else: libraries that print there (NCCL's version banner under NCCL_DEBUG=VERSION,
    torchrun notices) are sent to stderr for the whole run; emit() writes the line to the real stdout"""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-configuration lines (C1..C4, generator)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    import rsp_chains_b200 as R

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    warmup = max(3, args.warmup)

    # the job = world * FRAMES_PER_GPU frames, sharded in contiguous blocks (shard_rows); each frame has its own synthetic content
    f_lo, f_hi = R.shard_rows(world * FRAMES_PER_GPU, world, rank)
    assert f_hi - f_lo == FRAMES_PER_GPU
    rows_per_frame = N_RX * CHIRPS_PER_FRAME
    rows = rows_per_frame * FRAMES_PER_GPU
    samples_per_step = rows * N_POINTS
    reps = rows_per_frame // DISTINCT_ROWS
    bases = [host_rows(f) for f in range(f_lo, f_hi)]
    d_in = torch.cat([torch.from_numpy(b.view(np.int32)).cuda().repeat(reps, 1) for b in bases]).contiguous()   # [rows, N] resident in HBM
    d_out = torch.empty_like(d_in)
    assert d_in.shape[0] == rows

    ch = R.FftMagCfarChainVanilla(chain_params(), device=local_rank)
    ch.configure(LOG_N, 2, REGS)
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_chirps = CHIRPS_PER_FRAME * FRAMES_PER_GPU
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    barrier()
    for _ in range(warmup):
        ch.run(d_in, n_chirps, N_RX, out=d_out)
    barrier()
    sampler.mark()
    launches0 = ch.kernel_launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    for i in range(args.steps):
        ch.run(d_in, n_chirps, N_RX, out=d_out)
        ev[i + 1].record(stream)
    barrier()
    launches = ch.kernel_launches - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop() if rank == 0 else None
    max_ms = max_over_ranks(total_ms)
    value = world * samples_per_step * args.steps / (max_ms * 1e-3) / 1e6

    # e2e: the same step through rsp_run_host, pinned host buffers, H2D + D2H inside the timed region
    h_in = R.host_alloc(rows * N_POINTS * 4)
    h_out = R.host_alloc(rows * N_POINTS * 4)
    hv = h_in.reshape(rows, N_POINTS)
    for fi, b in enumerate(bases):
        for r in range(reps):
            r0 = fi * rows_per_frame + r * DISTINCT_ROWS
            hv[r0:r0 + DISTINCT_ROWS] = b
    for _ in range(2):
        ch.run_host_raw(h_in.ctypes.data, n_chirps, N_RX, h_out.ctypes.data)
    barrier()
    l0 = ch.kernel_launches
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        ch.run_host_raw(h_in.ctypes.data, n_chirps, N_RX, h_out.ctypes.data)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_launches = ch.kernel_launches - l0
    e2e_value = world * samples_per_step * args.e2e_steps / e2e_s / 1e6
    host_equal = bool((h_out.reshape(rows, N_POINTS)[:8] == d_out[:8].cpu().numpy().view(np.uint32)).all())

    # e2e, detections only: h_out = NULL, the detection list is the only result that crosses PCIe back (rsp_run_host with h_dets)
    max_dets = 1 << 22
    n_dets = 0
    for _ in range(1):
        n_dets = ch.run_host_dets_raw(h_in.ctypes.data, n_chirps, N_RX, max_dets)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        n_dets = ch.run_host_dets_raw(h_in.ctypes.data, n_chirps, N_RX, max_dets)
    torch.cuda.synchronize()
    e2e_det_s = max_over_ranks(time.perf_counter() - t0)
    e2e_det_value = world * samples_per_step * args.e2e_steps / e2e_det_s / 1e6
    n_peaks_out = int((h_out.reshape(rows, N_POINTS) & 1).sum())     # peak bits of the full-output run: must equal the list length

    # detections: the only inter-GPU exchange of the path; timed on its own (one frame's list per rank, all-gather of counts + padded lists)
    det_list = torch.zeros((max_dets, 4), dtype=torch.int32, device="cuda")
    det_count = torch.zeros(1, dtype=torch.int32, device="cuda")
    ch.run(d_in, n_chirps, N_RX, out=d_out, det_list=det_list, det_count=det_count)
    torch.cuda.synchronize()
    slot = 1 << 18
    gather_api = "rsp_gather_detections (C ABI: ncclAllGather of counts + one slot of %d entries per rank)" % slot
    try:
        comm = R.NcclComm(local_rank)
        R.gather_detections_nccl(ch, comm, det_list, det_count, slot)      # warm-up
        barrier()
        t0 = time.perf_counter()
        gathered, counts = R.gather_detections_nccl(ch, comm, det_list, det_count, slot)
        torch.cuda.synchronize()
        gather_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        n_gathered = int(torch.clamp(counts, max=slot).sum().item())
        comm.close()
    except Exception as e:      # no NCCL library on the box: the torch.distributed helper does the same exchange
        gather_api = f"gather_detections (torch.distributed all_gather; the C ABI path failed: {e!r})"
        gathered = R.gather_detections(det_list, det_count)
        barrier()
        t0 = time.perf_counter()
        gathered = R.gather_detections(det_list, det_count)
        torch.cuda.synchronize()
        gather_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        n_gathered = int(gathered.shape[0])

    if rank == 0:
        peak, peak_src = measured_peak()
        med_ms = float(np.median(per_launch_ms))
        avg_ms = total_ms / args.steps
        achieved = samples_per_step * BYTES_PER_SAMPLE / (avg_ms * 1e-3) / 1e9
        traffic = ncu_traffic() or (None, None)
        out = {
            "metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int16 fixed point (Q16.0, int32 arithmetic)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "variant": ch.variant, "samples_per_step_per_gpu": samples_per_step,
                       "l2_policy": "inputs+outputs (4 GiB per step) exceed the 126 MB L2",
                       "parallelism": f"{world * FRAMES_PER_GPU} frame(s) sharded over {world} GPU(s) (shard_rows), no data-path collective",
                       "detections_gathered": n_gathered, "host_path_equals_device_path": host_equal},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic[0], "traffic_source": traffic[1],
                         "algorithmic_bytes_per_launch": samples_per_step * BYTES_PER_SAMPLE, "peak_source": peak_src, "kernel_ms_avg": avg_ms, "kernel_ms_median": med_ms,
                         "bytes_per_sample": BYTES_PER_SAMPLE},
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": rows * N_POINTS * 4, "d2h_bytes_per_step": rows * N_POINTS * 4,
                    "steps": args.e2e_steps, "api": "rsp_run_host (pinned host buffers, 3-deep chunk pipeline)",
                    "detections_only": {"value": e2e_det_value, "unit": "Msamples/s", "h2d_bytes_per_step": rows * N_POINTS * 4,
                                        "d2h_bytes_per_step": int(min(n_dets, max_dets)) * 16 + 4, "detections_per_step": int(n_dets),
                                        "peak_bits_in_full_output": n_peaks_out,
                                        "api": "rsp_run_host(h_out = NULL, h_dets): the detection list is the only result copied back"}},
            "gather_ms": gather_ms, "gather": f"{gather_api}: {n_gathered} detections x 16 B over {world} rank(s), outside the timed step",
            "gpu_launches": int(launches), "e2e_gpu_launches": int(e2e_launches), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = time_oracle(bases[0])
            out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port", "sample": sample}
        else:
            out["cpu_baseline"] = None
    R.host_free(h_in); R.host_free(h_out)
    del d_in, d_out, det_list
    torch.cuda.empty_cache()
    if rank == 0:
        if world == 1 and not args.no_configs:
            try:
                out["configs"] = config_lines(torch, peak)
            except Exception as e:   # the headline line must survive a failing side configuration
                out["configs"] = [{"error": repr(e)}]
        emit(out)
    ch.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
