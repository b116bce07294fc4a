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
