"""Range-Doppler processing out of the reference's blocks.  The reference stops at the 1-D range CFAR; a radar user's
next step is the slow-time FFT over the chirps of a frame, the combining of the RX channels and a CFAR on the range-Doppler map.  All of
it is the reference's own arithmetic -- the SDF-FFT generator instantiated a second time with numPoints = chirps per frame, LogMagMux,
the CFAR cores -- plus two data movements (corner turn, RX mean) that are this library's (rsp_corner_turn, rsp_mag_combine_run).

    frame [rx][chirp][sample]  --range FFT-->  [rx][chirp][range]  --corner turn-->  [rx][range][chirp]  --Doppler FFT-->
    [rx][range][doppler]  --LogMagMux + mean over rx-->  [range][doppler]  --CFAR along Doppler-->  out [range][doppler]
    or, cfar_axis = "range":  --corner turn-->  [doppler][range]  --CFAR along range-->  out [doppler][range]
"""
from __future__ import annotations

from .chain import FftMagCfarChainVanilla
from .params import FftMagCfarVanillaParameters


class RangeDopplerChain:
    """`range_params`: the fast-time chain (its FFT block is used; its CFAR block when cfar_axis = "range").
    `doppler_params`: the slow-time chain, fftParams.numPoints = chirps per frame (FFT and LogMagMux blocks; CFAR when cfar_axis = "doppler").
    Registers are written on `.range` / `.doppler` as on any FftMagCfarChainVanilla."""

    def __init__(self, range_params: FftMagCfarVanillaParameters, doppler_params: FftMagCfarVanillaParameters, n_rx: int,
                 cfar_axis: str = "doppler", device: int = 0):
        assert cfar_axis in ("doppler", "range")
        assert n_rx >= 1 and n_rx & (n_rx - 1) == 0, "n_rx must be a power of two (the RX mean is a shift)"
        self.range = FftMagCfarChainVanilla(range_params, device=device)
        self.doppler = FftMagCfarChainVanilla(doppler_params, device=device)
        self.n_rx, self.cfar_axis = n_rx, cfar_axis
        self._scratch = None

    def close(self):
        self.range.close(); self.doppler.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def kernel_launches(self) -> int:
        return self.range.kernel_launches + self.doppler.kernel_launches

    def _buffers(self, like, cells: int):
        import torch
        if self._scratch is None or self._scratch[0].numel() < like.numel() or self._scratch[0].dtype != like.dtype:
            mag_shape = (cells,) if like.is_floating_point() else (cells,)
            self._scratch = (torch.empty_like(like), torch.empty_like(like),
                             torch.empty(mag_shape, dtype=torch.float32 if like.is_floating_point() else like.dtype, device=like.device),
                             torch.empty(mag_shape, dtype=torch.float32 if like.is_floating_point() else like.dtype, device=like.device))
        return self._scratch

    def run(self, d_in, out=None, peak_mask=None, det_list=None, det_count=None, stream=None):
        """d_in: one frame, device tensor [n_rx, chirps, N] of beats (fp32 flavour: [n_rx, chirps, N, 2]); out / peak_mask / det_list as
        for FftMagCfarChainVanilla.cfar, rows = range bins (cfar_axis "doppler") or Doppler bins ("range")."""
        n, m, n_rx = self.range.fftSize, self.doppler.fftSize, self.n_rx
        assert d_in.shape[0] == n_rx and d_in.shape[1] == m and d_in.shape[2] == n, (tuple(d_in.shape), (n_rx, m, n))
        a, b, mag, mag_t = self._buffers(d_in, n * m)
        self.range.fft(d_in, n_rx * m, a, stream=stream)                          # [rx][chirp][range]
        self.range.corner_turn(a, m, n, b, batch=n_rx, stream=stream)             # [rx][range][chirp]
        self.doppler.fft(b, n_rx * n, a, stream=stream)                           # [rx][range][doppler]
        self.doppler.mag_combine(a, n * m, n_rx, mag, stream=stream)              # [range][doppler]
        if self.cfar_axis == "doppler":
            return self.doppler.cfar(mag, n, out=out, peak_mask=peak_mask, det_list=det_list, det_count=det_count, stream=stream)
        self.range.corner_turn(mag, n, m, mag_t, stream=stream)                   # [doppler][range]
        return self.range.cfar(mag_t, m, out=out, peak_mask=peak_mask, det_list=det_list, det_count=det_count, stream=stream)

    def stages(self, d_in, stream=None):
        """the intermediate maps of `run` (copies), for tests: range FFT, Doppler FFT input, Doppler FFT, combined magnitude"""
        n, m, n_rx = self.range.fftSize, self.doppler.fftSize, self.n_rx
        a, b, mag, _ = self._buffers(d_in, n * m)
        out = {}
        self.range.fft(d_in, n_rx * m, a, stream=stream); out["range_fft"] = a.clone()
        self.range.corner_turn(a, m, n, b, batch=n_rx, stream=stream); out["turned"] = b.clone()
        self.doppler.fft(b, n_rx * n, a, stream=stream); out["doppler_fft"] = a.clone()
        self.doppler.mag_combine(a, n * m, n_rx, mag, stream=stream); out["combined"] = mag.clone()
        return out
