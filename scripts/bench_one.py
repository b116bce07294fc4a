#!/usr/bin/env python
"""One configuration of scripts/bench_configs.py by key, for profiling: python scripts/bench_one.py c3|c5f|c2|c4os|c5 [steps]"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import bench_configs as B
from rsp_chains_b200.params import GOSCFARType
import torch
key = sys.argv[1] if len(sys.argv) > 1 else "c3"
torch.cuda.set_device(0)
ca = lambda n, W, G, ds, mode=0: dict(fftWin=n, thresholdScaler=4, logOrLinearMode=1, divSum=ds, peakGrouping=0, cfarMode=mode, windowCells=W, guardCells=G)
if key == "c3":
    B.run_case("C3 x 16 frames fp32 GO", B.f32_params(2048, 32, 4), 11, ca(2048, 32, 4, 5, 1), 32768, 4, fp32=True, bytes_per_sample=12.125)
elif key == "c5f":
    B.run_case("C5 shape fp32 CA, 1/4 frame", B.f32_params(16384, 64, 8), 14, ca(16384, 64, 8, 6), 8192, 8, fp32=True, bytes_per_sample=12.125)
elif key == "c2":
    B.run_case("C2 x 256 frames", B.fixed_params(1024, 64, 4), 10, ca(1024, 16, 4, 4), 65536, 1)
elif key == "c5":
    B.run_case("C5 fixed one frame", B.fixed_params(16384, 64, 8), 14, ca(16384, 64, 8, 6), 32768, 8)
elif key in ("app12", "app12_14", "app12pg"):   # FftMagCfarChainVanillaApp / RxFftMagCfarTxChainApp: binPoint 12 (src/main/scala/FftMagCfarChain.scala:76-120)
    import rsp_chains_b200 as R
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    if key in ("app12", "app12pg"):
        B.run_case("FftMagCfarChainVanillaApp (binPoint 12) x 256 frames of 256 chirps", p, 10, dict(fftWin=1024, thresholdScaler=int(3.5 * 4096), logOrLinearMode=1, divSum=6, peakGrouping=int(key == "app12pg"), cfarMode=0, windowCells=64, guardCells=4), 65536, 1)
    else:
        p.fftParams.numPoints = 16384; p.cfarParams.fftSize = 16384
        p.fftParams.keepMSBorLSB = [True] * 14; p.fftParams.expandLogic = [0] * 14
        B.run_case("the same parameters at 16384 points", p, 14, dict(fftWin=16384, thresholdScaler=int(3.5 * 4096), logOrLinearMode=1, divSum=6, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4), 4096, 1)
elif key in ("c4cfar", "c4cfarb"):      # C4 "CFAR-only on stored 16-bit magnitude profiles": rsp_cfar_run, 1024 x 4096 x 16 frames
    from rsp_chains_b200.params import OSGoldenCFARType
    import rsp_chains_b200 as R, json, numpy as np
    n, rows = 4096, 16384
    b = key == "c4cfarb"
    regs = dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=32 if b else 64, guardCells=4, indexLagg=48, indexLead=48)
    rng = np.random.default_rng(1)
    prof = np.floor(rng.exponential(256.0, (64, n))).astype(np.int32); prof[:, [300, 1200, 2500, 3900]] += 4096
    d_in = torch.from_numpy(prof).cuda().repeat(rows // 64, 1).contiguous(); d_out = torch.empty_like(d_in)
    with R.FftMagCfarChainVanilla(B.fixed_params(n, 64, 4, OSGoldenCFARType if b else GOSCFARType)) as ch:
        ch.configure(12, 2, regs)
        for _ in range(3): ch.cfar(d_in, rows, out=d_out)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
        ev[0].record()
        for i in range(10):
            ch.cfar(d_in, rows, out=d_out); ev[i + 1].record()
        torch.cuda.synchronize()
        ms = float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(10)]))
        print(json.dumps({"config": key + " (CFAR block alone on stored profiles)", "rows": rows, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(rows * n / ms / 1e3, 1),
                          "GBps_algorithmic": round(rows * n * 8 / ms / 1e6, 1)}))
elif key == "c5pg":
    r = ca(16384, 64, 8, 6); r["peakGrouping"] = 1
    B.run_case("C5 fixed one frame, peakGrouping = 1", B.fixed_params(16384, 64, 8), 14, r, 32768, 8)
elif key == "c2pg":
    r = ca(1024, 16, 4, 4); r["peakGrouping"] = 1
    B.run_case("C2 x 256 frames, peakGrouping = 1", B.fixed_params(1024, 64, 4), 10, r, 65536, 1)
elif key == "c4os":
    os_regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
    B.run_case("C4 GOS", B.fixed_params(4096, 64, 4, GOSCFARType), 12, os_regs, 1024, 1)
elif key in ("c4os16", "c4os16pg"):
    os_regs = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=int(key == "c4os16pg"), cfarMode=0, windowCells=64, guardCells=4, indexLagg=48, indexLead=48)
    B.run_case("C4 x 16 GOS" + (", peakGrouping = 1" if key == "c4os16pg" else ""), B.fixed_params(4096, 64, 4, GOSCFARType), 12, os_regs, 16384, 1)
elif key in ("c4os16d", "c4osbd"):      # the same on full-range random beats: every plane is populated everywhere (worst case of the bit-plane selection)
    from rsp_chains_b200 import synth
    from rsp_chains_b200.params import OSGoldenCFARType
    import numpy as np
    n, rows = 4096, 16384
    b = key == "c4osbd"
    regs = dict(fftWin=n, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=32 if b else 64, guardCells=4, indexLagg=48, indexLead=48)
    import rsp_chains_b200 as R, json
    base = torch.from_numpy(synth.random_beats(64, n, seed=3).view(np.int32)).cuda()
    d_in = base.repeat(rows // 64, 1).contiguous(); d_out = torch.empty_like(d_in)
    with R.FftMagCfarChainVanilla(B.fixed_params(n, 64, 4, OSGoldenCFARType if b else GOSCFARType)) as ch:
        ch.configure(12, 2, regs)
        for _ in range(3): ch.run(d_in, rows, 1, out=d_out)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
        ev[0].record()
        for i in range(10):
            ch.run(d_in, rows, 1, out=d_out); ev[i + 1].record()
        torch.cuda.synchronize()
        ms = float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(10)]))
        print(json.dumps({"config": key + " (full-range random beats)", "variant": ch.variant, "rows": rows, "n": n, "ms": round(ms, 4), "Msamples_per_s": round(rows * n / ms / 1e3, 1)}))
elif key == "c4osb":
    from rsp_chains_b200.params import OSGoldenCFARType
    os_b = dict(fftWin=4096, thresholdScaler=2, logOrLinearMode=1, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    B.run_case("C4 x 16 OS golden (reading B)", B.fixed_params(4096, 64, 4, OSGoldenCFARType), 12, os_b, 16384, 1)
elif key.startswith("small"):   # smallS: fixed CA chain at N = 2^S < 1024 (run-time-parameter kernel, MODE_FAST instantiation)
    S = int(key[5:]); n = 1 << S
    B.run_case(f"fixed CA N={n}", B.fixed_params(n, 16, 2), S, ca(n, 16, 2, 4), (1 << 26) // n, 1)
elif key.startswith("fx"):      # fxS: fixed CA chain at N = 2^S, 512 MiB footprint
    S = int(key[2:]); n = 1 << S
    B.run_case(f"fixed CA N={n}", B.fixed_params(n, 64, 8), S, ca(n, 64, 8, 6), (1 << 26) // n, 1)
elif key.startswith("fp"):      # fpS: fp32 CA chain at N = 2^S
    S = int(key[2:]); n = 1 << S
    B.run_case(f"fp32 CA N={n}", B.f32_params(n, 64, 8), S, ca(n, 64, 8, 6), (1 << 25) // n, 1, fp32=True, bytes_per_sample=12.125)
