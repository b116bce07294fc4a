"""rsp_run_from_generator_static -- the chain fed by the PLFG / NCO generator with its scratch beats in a file-scope device array of the
library (no cudaMalloc in the call; no counterpart in the reference) -- against the existing
rsp_run_from_generator, through the C ABI, bit for bit (integer path: the tolerance of the other fixed-point tests is 0), on two calls in a
row with different schedules and sizes: the second call must not see the first one's beats or segment list.  In a file of its own that
pytest collects last, so that a failure here cannot hide another test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

import rsp_chains_b200 as R
from rsp_chains_b200 import _abi
from oracle.glue import write_plfg
from tests.util import configure, from_dev_u32

pytestmark = pytest.mark.gpu

CHIRPS_A = [[(3, 200, True, False), (5, 120, False, False)], [(1, 255, True, False), (0, 40, False, True), (2, 180, False, False)]]
CHIRPS_B = [[(7, 150, True, False)], [(11, 90, False, False), (11, 90, True, False), (4, 33, True, False)], [(2, 60, True, False)]]


def test_static_scratch_entry_point_equals_the_allocating_one(cuda, oracle):
    p = R.rsp_chain_vanilla_app_params()
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=5, peakGrouping=0, cfarMode=0, windowCells=32, guardCells=4)
    cases = [(CHIRPS_A, [1, 0, 1], [2, 1, 1], 37, 24, 1),      # (chirps, order, repeats, startingPoint, chirps to process, n_rx)
             (CHIRPS_B, [2, 0, 1, 0], [1, 3, 1, 2], 5, 7, 2)]  # second call: other schedule, fewer samples than the first
    with R.RspChainVanilla(p) as ch:
        configure(ch, 10, 2, regs)
        for chirps, order, repeats, start, n_chirps, n_rx in cases:
            write_plfg(ch, p, oracle, start, chirps, order, repeats, 0, 10)
            rows = n_chirps * n_rx
            want = cuda.zeros((rows, 1024), dtype=cuda.int32, device="cuda")
            got = cuda.zeros((rows, 1024), dtype=cuda.int32, device="cuda")
            dl_w, dc_w = cuda.zeros((rows * 1024, 4), dtype=cuda.int32, device="cuda"), cuda.zeros(1, dtype=cuda.int32, device="cuda")
            dl_g, dc_g = cuda.zeros((rows * 1024, 4), dtype=cuda.int32, device="cuda"), cuda.zeros(1, dtype=cuda.int32, device="cuda")
            ch.run_from_generator(n_chirps, n_rx, out=want, det_list=dl_w, det_count=dc_w)
            cuda.cuda.synchronize()
            launches = ch.kernel_launches
            ch.run_from_generator_static(n_chirps, n_rx, out=got, det_list=dl_g, det_count=dc_g)
            cuda.cuda.synchronize()
            assert ch.kernel_launches - launches == 2                      # generator_kernel + the chain kernel
            w, g = from_dev_u32(want), from_dev_u32(got)
            assert (w & 1).sum() > 0                                        # the tone is detected: the comparison is not of two empty outputs
            assert (g == w).all(), f"first difference at {np.argwhere(g != w)[:3].tolist()}"
            k = int(dc_w.item())
            assert int(dc_g.item()) == k
            assert sorted(map(tuple, dl_g[:k].cpu().numpy().tolist())) == sorted(map(tuple, dl_w[:k].cpu().numpy().tolist()))
        # beyond the static scratch: refused with the limit in the message, nothing launched
        launches = ch.kernel_launches
        too_many = (1 << 20) // 1024 + 1
        out = cuda.zeros((too_many, 1024), dtype=cuda.int32, device="cuda")
        st = _abi.lib().rsp_run_from_generator_static(ch._h, too_many, 1, out.data_ptr(), None, None, None)
        assert st == _abi.RSP_EINVAL and "RSP_STATIC_SCRATCH_BEATS" in _abi.lib().rsp_last_error().decode()
        assert ch.kernel_launches == launches
        # exactly the limit is accepted
        ch.run_from_generator_static(too_many - 1, 1, out=out)
        ref = cuda.zeros_like(out)
        ch.run_from_generator(too_many - 1, 1, out=ref)
        cuda.cuda.synchronize()
        assert bool((out[:too_many - 1] == ref[:too_many - 1]).all().item())


def test_static_entry_point_without_a_generator_is_refused(cuda):
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    with R.FftMagCfarChainVanilla(p) as ch:
        out = cuda.zeros((1, 1024), dtype=cuda.int32, device="cuda")
        st = _abi.lib().rsp_run_from_generator_static(ch._h, 1, 1, out.data_ptr(), None, None, None)
        assert st == _abi.RSP_EINVAL
