#!/usr/bin/env python
"""bench.py -- the headline benchmark: end-to-end chain Msamples/s (FFT -> LogMagMux -> CFAR), fixed point.

A step = one pass of the fused chain over one batch of synthetic frames.  Workload = BASELINE.json configs[1]
(one frame = 256 chirps x 1024-point fixed-point range FFT + JPL magnitude + CA-CFAR, guard 4, ref 16), batched as
FRAMES frames per step per GPU so that input + output (2 x 256 MiB) exceed the 126 MB L2.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  torchrun ... bench.py --gpus N ...      one rank per GPU, frames sharded (weak scaling), NCCL only gathers detections

Prints ONE JSON line (rank 0).  `value` is device-resident throughput, `e2e` goes through rsp_run_host with pinned
host buffers (H2D and D2H inside the timed region), `roofline` relates the fused kernel to the measured HBM peak,
`cpu_baseline` times the CPU oracle (a port of the reference's arithmetic; the Scala/Chisel reference itself cannot run
here) on a bounded sample.
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

N_POINTS = 1024
CHIRPS_PER_FRAME = 256
FRAMES = 256                      # frames per step per GPU -> 65536 chirps = 64 Mi samples, 256 MiB in + 256 MiB out
DISTINCT_FRAMES = 16              # synthesised on the host, tiled on the device
BYTES_PER_SAMPLE = 8              # Cat(re16, im16) in + Cat(threshold, bin, peak) out  
REGS = dict(fftWin=N_POINTS, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
WORKLOAD = (f"{FRAMES} frames x {CHIRPS_PER_FRAME} chirps x {N_POINTS}-pt fixed-point (Q16.0) range FFT + JPL mag + "
            f"CA-CFAR (guard 4, ref 16) per GPU per step [BASELINE configs[1] batched over frames]")


def chain_params():
    import rsp_chains_b200 as R
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    p.fftParams.binPoint = 0
    p.magParams.binPoint = 0
    p.cfarParams.protoIn = p.cfarParams.protoThreshold = p.cfarParams.protoScaler = R.FixedPoint(16, 0)
    return p


def ncu_traffic():
    """DRAM bytes per launch of the chain kernel from the committed `ncu --set full` capture (profiles/), or None"""
    import glob
    best = None
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_full_chain_fixed_fast*.txt"))):
        rd = wr = None
        for line in open(f):
            c = line.rstrip("\n").split("\t")
            if len(c) == 3 and c[0] == "dram__bytes_read.sum" and c[1] == "Mbyte":
                rd = float(c[2])
            if len(c) == 3 and c[0] == "dram__bytes_write.sum" and c[1] == "Mbyte":
                wr = float(c[2])
        if rd is not None and wr is not None:
            best = ((rd + wr) * 1e6, os.path.basename(f))
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
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
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
        for r in self.rows:
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
    fft, mag, cfar = oracle_cfgs(O, chain_params(), 10, 2, REGS)
    return O, fft, mag, cfar


def time_oracle(beats, target_s=12.0):
    """all host threads, bounded sample sized to ~target_s of CPU work"""
    O, fft, mag, cfar = oracle_setup()
    threads = host_threads()
    probe = beats[:256]
    t0 = time.perf_counter(); O.chain_fixed_words(fft, mag, cfar, probe, n_threads=threads); dt = time.perf_counter() - t0
    rate = probe.size / max(dt, 1e-9)
    rows = int(max(256, (rate * target_s) // N_POINTS))
    reps = max(1, -(-rows // beats.shape[0]))
    sample = beats[:min(rows, beats.shape[0])]
    t0 = time.perf_counter()
    for _ in range(reps):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    dt = time.perf_counter() - t0
    return (sample.size * reps / dt / 1e6, threads,
            f"{reps} x {sample.shape[0]} chirps x {N_POINTS} of the same synthetic workload, {dt:.1f} s wall, OpenMP over chirps")


def host_frames():
    from rsp_chains_b200 import synth
    return synth.chirps_fixed(DISTINCT_FRAMES * CHIRPS_PER_FRAME, N_POINTS, seed=11110)


def run_reference(args):
    """the reference arm: the CPU implementation of the path (oracle port; the Scala/Chisel reference cannot run here)"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O, fft, mag, cfar = oracle_setup()
    threads = host_threads()
    base = host_frames()
    # each step: a bounded sample of the workload (~1.5 s of CPU work)
    t0 = time.perf_counter(); O.chain_fixed_words(fft, mag, cfar, base[:256], n_threads=threads); dt = time.perf_counter() - t0
    rows = int(min(base.shape[0], max(256, (256 * 1.5 / max(dt, 1e-9)))))
    sample = base[:rows]
    for _ in range(args.warmup):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.chain_fixed_words(fft, mag, cfar, sample, n_threads=threads)
    dt = time.perf_counter() - t0
    val = sample.size * args.steps / dt / 1e6
    print(json.dumps({
        "impl": "reference", "metric": "end-to-end chain Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int16 fixed point (Q16.0, int64 intermediates)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{rows} chirps x {N_POINTS} per step on the host"},
        "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": threads, "kind": "port",
                         "sample": f"{rows} chirps x {N_POINTS} per step, {args.steps} steps"},
        "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is a Scala/Chisel hardware generator (no JVM here): CPU arm = oracle/ C port of its arithmetic, OpenMP over chirps",
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    import rsp_chains_b200 as R
    from rsp_chains_b200 import _abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    warmup = max(3, args.warmup)

    rows = FRAMES * CHIRPS_PER_FRAME
    samples_per_step = rows * N_POINTS
    base = host_frames()
    d_base = torch.from_numpy(base.view(np.int32)).cuda()
    d_in = d_base.repeat(FRAMES // DISTINCT_FRAMES, 1).contiguous()       # [rows, N] resident in HBM
    d_out = torch.empty_like(d_in)
    assert d_in.shape[0] == rows

    ch = R.FftMagCfarChainVanilla(chain_params(), device=local_rank)
    ch.configure(10, 2, REGS)
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        ch.run(d_in, rows, 1, out=d_out)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ch.kernel_launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    for i in range(args.steps):
        ch.run(d_in, rows, 1, out=d_out)
        ev[i + 1].record(stream)
    barrier()
    launches = ch.kernel_launches - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_ms = float(t.item())
    value = world * samples_per_step * args.steps / (max_ms * 1e-3) / 1e6

    # parity guard on the bench data itself: first frame against the oracle (rank 0, outside the timed region)
    # e2e: the same step through rsp_run_host, pinned host buffers, H2D + D2H inside the timed region
    h_in = R.host_alloc(rows * N_POINTS * 4)
    h_out = R.host_alloc(rows * N_POINTS * 4)
    h_in.reshape(rows, N_POINTS)[:] = np.tile(base, (FRAMES // DISTINCT_FRAMES, 1))
    for _ in range(2):
        ch.run_host_raw(h_in.ctypes.data, rows, 1, h_out.ctypes.data)
    barrier()
    l0 = ch.kernel_launches
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        ch.run_host_raw(h_in.ctypes.data, rows, 1, h_out.ctypes.data)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_launches = ch.kernel_launches - l0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * samples_per_step * args.e2e_steps / float(t.item()) / 1e6
    host_equal = bool((h_out.reshape(rows, N_POINTS)[:64] == d_out[:64].cpu().numpy().view(np.uint32)).all())

    # detections: the only inter-GPU exchange of the path (outside the timed region)
    det_list = torch.zeros((1 << 20, 4), dtype=torch.int32, device="cuda")
    det_count = torch.zeros(1, dtype=torch.int32, device="cuda")
    ch.run(d_in[:CHIRPS_PER_FRAME], CHIRPS_PER_FRAME, 1, out=d_out[:CHIRPS_PER_FRAME], det_list=det_list, det_count=det_count)
    torch.cuda.synchronize()
    gathered = R.gather_detections(det_list, det_count)
    n_gathered = int(gathered.shape[0])

    if rank == 0:
        peak, peak_src = measured_peak()
        med_ms = float(np.median(per_launch_ms))
        avg_ms = total_ms / args.steps
        achieved = samples_per_step * BYTES_PER_SAMPLE / (avg_ms * 1e-3) / 1e9
        out = {
            "metric": "end-to-end chain Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int16 fixed point (Q16.0, int32 arithmetic)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "variant": ch.variant, "samples_per_step_per_gpu": samples_per_step,
                       "l2_policy": "inputs+outputs (512 MiB per step) exceed the 126 MB L2", "parallelism": f"frames sharded over {world} GPU(s), no data-path collective",
                       "detections_gathered_one_frame": n_gathered, "host_path_equals_device_path": host_equal},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (ncu_traffic() or (None, None))[0], "traffic_source": (ncu_traffic() or (None, None))[1],
                         "algorithmic_bytes_per_launch": samples_per_step * BYTES_PER_SAMPLE, "peak_source": peak_src, "kernel_ms_avg": avg_ms, "kernel_ms_median": med_ms,
                         "bytes_per_sample": BYTES_PER_SAMPLE},
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": rows * N_POINTS * 4, "d2h_bytes_per_step": rows * N_POINTS * 4,
                    "steps": args.e2e_steps, "api": "rsp_run_host (pinned host buffers, 3-deep chunk pipeline)"},
            "gpu_launches": int(launches), "e2e_gpu_launches": int(e2e_launches), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = time_oracle(base)
            out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port", "sample": sample}
        else:
            out["cpu_baseline"] = None
        print(json.dumps(out))
    R.host_free(h_in); R.host_free(h_out)
    ch.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
