#!/usr/bin/env python
"""Throughput of configurations that the specialised kernels do not take (the run-time-parameter kernel): where are the cliffs?
python scripts/cliff_probe.py   (one JSON line per case; 4096 rows x 1024 points)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import bench_configs as B
from rsp_chains_b200.params import CACFARType, GOSCACFARType, GOSCFARType
import torch
torch.cuda.set_device(0)
base = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4, indexLagg=12, indexLead=12, subWindowSize=4)
rows = 4096
def case(name, p, regs, **kw):
    try:
        B.run_case(name, p, 10, regs, rows, 1, **kw)
    except Exception as e:      # noqa: BLE001
        print({"config": name, "error": str(e)[:200]})
case("CA (specialised, for scale)", B.fixed_params(1024, 64, 4), base)
case("GOSCA type, cfarAlgorithm = 1 (GOS)", B.fixed_params(1024, 64, 4, GOSCACFARType), dict(base, cfarAlgorithm=1))
case("GOSCA type, cfarAlgorithm = 0 (CA on the Lis core)", B.fixed_params(1024, 64, 4, GOSCACFARType), dict(base, cfarAlgorithm=0))
p = B.fixed_params(1024, 64, 4); p.cfarParams.includeCASH = True; p.cfarParams.minSubWindowSize = 4
case("CASH (ASR core), cfarMode 3", p, dict(base, cfarMode=3, divSum=2))
p = B.fixed_params(1024, 64, 4); p.fftParams.trimType = "Convergent"
case("CA, Convergent rounding", p, base)
import rsp_chains_b200 as R
p = R.fft_mag_cfar_chain_vanilla_app_params(); p.fftParams.trimType = "Convergent"
case("FftMagCfarChainVanillaApp (binPoint 12), Convergent", p, dict(base, thresholdScaler=int(3.5 * 4096), divSum=6, windowCells=64))
p = B.fixed_params(1024, 64, 4); p.fftParams.trimType = "RoundDown"
case("CA, RoundDown (any other trim type)", p, base)
p = B.fixed_params(1024, 64, 4); p.cfarParams.sendCut = True
case("CA, sendCut", p, base)
