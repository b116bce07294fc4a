"""Synthetic FMCW beat signals for tests and benchmarks (host-side numpy; not part of the hot path).

Recipe adapted from the reference's own test signals (src/test/scala/RspChainTesterUtils.scala:56-91,
src/main/scala/MemForTesting.scala:27-40): complex tones (targets) plus noise, quantised to 16-bit beats
Cat(re16, im16).  Amplitudes keep the chain away from the overflow corners listed in DESIGN.md."""
from __future__ import annotations

import numpy as np


def pack_iq(re, im) -> np.ndarray:
    re = np.asarray(re).astype(np.int64)
    im = np.asarray(im).astype(np.int64)
    return (((re & 0xFFFF) << 16) | (im & 0xFFFF)).astype(np.uint32)


def unpack_iq(words):
    w = np.asarray(words, dtype=np.uint32)
    re = (w >> 16).astype(np.uint16).view(np.int16).astype(np.int32)
    im = (w & 0xFFFF).astype(np.uint16).view(np.int16).astype(np.int32)
    return re, im


def chirps_complex(n_rows: int, n: int, seed: int = 11110, targets=((0.125, 0.4), (0.25, 0.2)), noise: float = 2.0 ** -7,
                   drift_rows: int = 16, rx_phase=None) -> np.ndarray:
    """complex128 [n_rows, n], unit full scale: targets at fractional frequencies drifting +1 bin every drift_rows rows,
    independent complex Gaussian noise."""
    rng = np.random.default_rng(seed)
    t = np.arange(n)[None, :]
    rows = np.arange(n_rows)[:, None]
    x = np.zeros((n_rows, n), np.complex128)
    for f, a in targets:
        binf = np.round(f * n) + (rows // drift_rows)
        ph = 2.0 * np.pi * (binf / n) * t
        if rx_phase is not None:
            ph = ph + np.asarray(rx_phase)[:, None]
        x += a * np.exp(1j * ph)
    x += noise * (rng.standard_normal((n_rows, n)) + 1j * rng.standard_normal((n_rows, n)))
    return x


def chirps_fixed(n_rows: int, n: int, seed: int = 11110, frac_bits: int = 13, **kw) -> np.ndarray:
    """uint32 beats [n_rows, n]: chirps_complex scaled by 2^frac_bits, floored, clipped to 16 bits"""
    x = chirps_complex(n_rows, n, seed, **kw) * (1 << frac_bits)
    re = np.clip(np.floor(x.real), -32767, 32767).astype(np.int32)
    im = np.clip(np.floor(x.imag), -32767, 32767).astype(np.int32)
    return pack_iq(re, im)


def chirps_f32(n_rows: int, n: int, seed: int = 11110, **kw) -> np.ndarray:
    """float32 [n_rows, n, 2]"""
    x = chirps_complex(n_rows, n, seed, **kw)
    return np.stack([x.real, x.imag], axis=-1).astype(np.float32)


def random_beats(n_rows: int, n: int, seed: int = 1, full_range: bool = True) -> np.ndarray:
    """uniform random 16-bit halves, including the -32768 / +32767 corners when full_range"""
    rng = np.random.default_rng(seed)
    if full_range:
        return rng.integers(0, 1 << 32, size=(n_rows, n), dtype=np.uint64).astype(np.uint32)
    re = rng.integers(-8192, 8192, size=(n_rows, n))
    im = rng.integers(-8192, 8192, size=(n_rows, n))
    return pack_iq(re, im)
