"""Glue between the product's host-side parameter classes and the CPU oracle's configs (tests only)."""
from __future__ import annotations

import numpy as np

from rsp_chains_b200.params import log2Up

from oracle.glue import oracle_cfgs  # noqa: F401


def configure(chain, log_n, sel, cfar_regs: dict):
    """write the run-time registers through the bus exactly as the reference testers do"""
    chain.configure(log_n, sel, cfar_regs)


def to_dev(torch, a: np.ndarray):
    if a.dtype == np.uint32:
        return torch.from_numpy(a.view(np.int32)).cuda()
    return torch.from_numpy(a).cuda()


def from_dev_u32(t) -> np.ndarray:
    return t.cpu().numpy().view(np.uint32)


def decode_cfar_words(words, fft_size):
    """threshold = out >> (log2 N + 1), bin, peak = out & 1 (src/test/scala/FftMagCfarChainTester.scala:163-167)"""
    w = np.asarray(words, dtype=np.uint32)
    b = log2Up(fft_size)
    return (w >> (b + 1)), (w >> 1) & ((1 << b) - 1), w & 1


NCO_FLAVOURS = [dict(), dict(phaseAccEnable=False), dict(phaseWidth=14, tableSize=1024), dict(phaseWidth=14, tableSize=256, tableWidth=12),
                dict(phaseWidth=12, tableSize=256), dict(rasterizedMode=True, tableSize=100), dict(rasterizedMode=True, tableSize=384)]


def random_plfg_programme(rng):
    """one random PLFG programme within RspChainVanillaApp's PLFGParams: (nco keywords, chirps, order, repeats, frameNum, pause, startingPoint,
    samples to ask of an endless schedule).  Shared by the CPU (emulated) and GPU randomised generator tests."""
    nco_kw = NCO_FLAVOURS[int(rng.integers(0, len(NCO_FLAVOURS)))]
    n_types = int(rng.integers(1, 5))
    chirps = [[(int(rng.integers(0, 1 << int(rng.integers(1, 12)))), int(rng.integers(0, 256)), bool(rng.integers(0, 2)), bool(rng.random() < 0.15))
               for _ in range(int(rng.integers(1, 5)))] for _ in range(n_types)]
    n_order = int(rng.integers(1, 9))
    order = [int(rng.integers(0, n_types)) for _ in range(n_order)]
    repeats = [int(rng.integers(0, 5)) for _ in range(n_order)]
    frame_num, inter, start = int(rng.integers(0, 4)), int(rng.integers(0, 64)), int(rng.integers(0, 1 << 16))
    return nco_kw, chirps, order, repeats, frame_num, inter, start, int(rng.integers(3000, 40000))
