#!/usr/bin/env python
"""e2e (host buffers, rsp_run_host) throughput of the bench.py workload (one C5 frame: 2 GiB in + 2 GiB out) for one setting of the
pipeline knobs, which are read once per process: RSP_PIPE_DEPTH=3 RSP_PIPE_CHUNK_MIB=16 python scripts/e2e_sweep.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import rsp_chains_b200 as R
import bench as B
torch.cuda.set_device(0)
rows = B.N_RX * B.CHIRPS_PER_FRAME
base = B.host_rows(0)
h_in = R.host_alloc(rows * B.N_POINTS * 4); h_out = R.host_alloc(rows * B.N_POINTS * 4)
h_in.reshape(rows, B.N_POINTS)[:] = np.tile(base, (rows // B.DISTINCT_ROWS, 1))
ch = R.FftMagCfarChainVanilla(B.chain_params(), device=0); ch.configure(B.LOG_N, 2, B.REGS)
for _ in range(2): ch.run_host_raw(h_in.ctypes.data, B.CHIRPS_PER_FRAME, B.N_RX, h_out.ctypes.data)
torch.cuda.synchronize(); t0 = time.perf_counter()
K = 4
for _ in range(K): ch.run_host_raw(h_in.ctypes.data, B.CHIRPS_PER_FRAME, B.N_RX, h_out.ctypes.data)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
gs = rows * B.N_POINTS * K / dt / 1e9
print(f"depth={os.environ.get('RSP_PIPE_DEPTH','3')} chunk={os.environ.get('RSP_PIPE_CHUNK_MIB','16')}MiB ramp={os.environ.get('RSP_PIPE_RAMP','1')}: {gs:.2f} Gsamples/s = {gs*4:.1f} GB/s each way")
