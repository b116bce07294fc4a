"""Multi-GPU path on real devices: rows sharded over two ranks with no data-path collective, detection lists gathered through the
C ABI's rsp_gather_detections (NCCL bound at run time) and, for comparison, through the torch.distributed helper.  Needs two GPUs on
one box (`gpurun --gpus 2`); the host-side logic of the same path is covered on CPU by tests/test_abi_host.py (gloo, world size 2)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, result_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import rsp_chains_b200 as R
    from rsp_chains_b200 import synth
    from oracle import oracle as O
    from oracle.glue import oracle_cfgs
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    p = R.fft_mag_cfar_chain_vanilla_app_params()
    p.fftParams.binPoint = 0
    p.magParams.binPoint = 0
    p.cfarParams.protoIn = p.cfarParams.protoThreshold = p.cfarParams.protoScaler = R.FixedPoint(16, 0)
    regs = dict(fftWin=1024, thresholdScaler=4, logOrLinearMode=1, divSum=4, peakGrouping=0, cfarMode=0, windowCells=16, guardCells=4)
    rows = 300
    beats = synth.chirps_fixed(rows, 1024, seed=123)        # the whole job, identical on every rank; each rank runs its block of rows
    lo, hi = R.shard_rows(rows, world, rank)
    with R.FftMagCfarChainVanilla(p, device=rank) as ch:
        ch.configure(10, 2, regs)
        d_in = torch.from_numpy(beats[lo:hi].view(np.int32)).cuda()
        d_out = torch.empty_like(d_in)
        det_list = torch.zeros((4096, 4), dtype=torch.int32, device="cuda")
        det_count = torch.zeros(1, dtype=torch.int32, device="cuda")
        ch.run(d_in, hi - lo, 1, out=d_out, det_list=det_list, det_count=det_count)
        comm = R.NcclComm(rank)
        gathered, counts = R.gather_detections_nccl(ch, comm, det_list, det_count, 2048)
        torch.cuda.synchronize()
        via_torch = R.gather_detections(det_list, det_count)
        comm.close()
    counts = counts.cpu().numpy()
    got = set()
    for r in range(world):
        r_lo, _ = R.shard_rows(rows, world, r)
        for d in gathered[r, :counts[r]].cpu().numpy():
            got.add((int(d[0]) + r_lo, int(d[1])))
    t_set = set()
    for d in via_torch.cpu().numpy():
        t_set.add((int(d[1]) + R.shard_rows(rows, world, int(d[0]))[0], int(d[2])))
    if rank == 0:
        fft, mag, cfar = oracle_cfgs(O, p, 10, 2, regs)
        want = O.chain_fixed_words(fft, mag, cfar, beats)
        exp = set(map(tuple, np.argwhere((want & 1) == 1).tolist()))
        ok = (got == exp) and (t_set == exp) and int(counts.sum()) == len(exp) and len(exp) > 500
        open(result_path, "w").write("ok" if ok else f"mismatch: {len(got)} {len(t_set)} {len(exp)} {counts.tolist()}")
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_rows_and_nccl_gather_two_gpus(cuda, tmp_path):
    if cuda.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one box (run with `gpurun --gpus 2`); single-GPU boxes cover the world-size-1 path in bench.py")
    import torch.multiprocessing as mp
    res = str(tmp_path / "result.txt")
    mp.spawn(_worker, args=(2, 29000 + os.getpid() % 1000, res), nprocs=2, join=True)
    assert open(res).read() == "ok"
