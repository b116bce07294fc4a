"""CPU-side checks of the drop-in boundary and the host logic (no compute calls: there is no GPU here).
 - librsp_b200.so loads and exports every symbol include/rsp_chains.h declares
 - the ctypes structs match the C structs (sizes through rsp_chain_params_default round trip)
 - parameter validation reproduces the reference's `require`s (status codes, messages)
 - without a CUDA device rsp_chain_create fails loudly (RSP_ECUDA): there is no CPU fallback
 - the product package never imports the oracle
 - detection gather + frame sharding under torch.distributed gloo, world_size 2"""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "rsp_chains.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rsp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from rsp_chains_b200 import _abi
    lib = _abi.lib()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"librsp_b200.so lacks {s}"
    assert sorted(_abi.EXPORTS) == syms
    assert lib.rsp_abi_version() == 3
    assert lib.rsp_status_str(1) == b"RSP_EINVAL"


def test_default_params_are_the_vanilla_app(tmp_path):
    """rsp_chain_params_default = RspChainVanillaApp (src/main/scala/RspChain.scala:83-147)"""
    from rsp_chains_b200 import _abi
    p = _abi.ChainParams()
    _abi.lib().rsp_chain_params_default(C.byref(p))
    assert (p.fft.numPoints, p.fft.dataWidth, p.fft.binPoint, p.fft.twiddleWidth) == (1024, 16, 0, 16)
    assert (p.nco.tableSize, p.nco.tableWidth, p.nco.phaseWidth) == (128, 16, 9)
    assert (p.cfar.leadLaggWindowSize, p.cfar.guardWindowSize, p.cfar.fftSize) == (64, 4, 1024)
    assert p.hasGenerator == 1 and p.plfg.maxNumOfSegments == 4 and p.mag.log2LookUpWidth == 9
    # struct layout: a C program sees the same sizeof
    src = tmp_path / "sz.c"
    src.write_text('#include "rsp_chains.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu", sizeof(rsp_chain_params), sizeof(rsp_detection), sizeof(rsp_dets));return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    assert [int(v) for v in out] == [C.sizeof(_abi.ChainParams), C.sizeof(_abi.Detection), C.sizeof(_abi.Dets)]


def _create(p):
    from rsp_chains_b200 import _abi
    h = C.c_void_p()
    st = _abi.lib().rsp_chain_create(C.byref(p), 0, C.byref(h))
    msg = _abi.lib().rsp_last_error().decode()
    if st == 0:
        _abi.lib().rsp_chain_destroy(h)
    return st, msg


def test_requires_and_no_cpu_fallback():
    import torch
    from rsp_chains_b200 import _abi
    base = _abi.ChainParams()
    _abi.lib().rsp_chain_params_default(C.byref(base))

    def variant(**kw):
        p = _abi.ChainParams.from_buffer_copy(base)
        for k, v in kw.items():
            obj, name = k.split("__")
            setattr(getattr(p, obj), name, v)
        return p

    st, msg = _create(variant(fft__numPoints=1000))
    assert st == _abi.RSP_EINVAL and "power of 2" in msg                       # SDFFFTParams.scala:53
    st, msg = _create(variant(cfar__guardWindowSize=0))
    assert st == _abi.RSP_EINVAL                                                # CFARParams.scala:41
    st, msg = _create(variant(cfar__leadLaggWindowSize=4, cfar__guardWindowSize=4))
    assert st == _abi.RSP_EINVAL                                                # CFARParams.scala:42
    st, msg = _create(variant(cfar__includeCASH=1, cfar__minSubWindowSize=0))
    assert st == _abi.RSP_EINVAL and "minSubWindowSize" in msg                  # CFARParams.scala:43-44
    st, msg = _create(variant(fft__expandLogic=1, fft__sdfRadix=2))             # grow stages: built for the radix 2^2 DIF chain only
    assert st == _abi.RSP_EUNSUPPORTED and "expandLogic" in msg
    st, msg = _create(variant(fft__dataWidth=24, fft__numPoints=16384))         # the wide FFT block keeps a row of 64-bit points in shared memory
    assert st == _abi.RSP_EUNSUPPORTED and "8192" in msg
    st, msg = _create(variant(fft__dataWidth=33))
    assert st == _abi.RSP_EUNSUPPORTED
    st, msg = _create(variant(fft__decimType=2))
    assert st == _abi.RSP_EINVAL
    st, msg = _create(variant(fft__sdfRadix=8))
    assert st == _abi.RSP_EINVAL and "Radix" in msg                             # SDFFFTParams.scala:61-64
    st, msg = _create(variant(nco__ditherEnable=1))
    assert st == _abi.RSP_EUNSUPPORTED
    st, msg = _create(variant(nco__tableSize=100))                               # NCOTable.scala:103
    assert st == _abi.RSP_EINVAL and "power of 2" in msg
    st, msg = _create(variant(nco__tableSize=100, nco__rasterizedMode=1, nco__phaseWidth=10))   # NCO.scala:56-62: not a parameter in this mode
    assert st == _abi.RSP_EINVAL and "log2Ceil(tableSize) + 2" in msg
    if not torch.cuda.is_available():
        st, msg = _create(base)
        assert st == _abi.RSP_ECUDA and "no CPU path" in msg


def test_params_mirror_reference_case_classes():
    import rsp_chains_b200 as R
    p = R.rsp_chain_vanilla_app_params()
    assert p.fftParams.numPoints == 1024 and p.fftParams.sdfRadix == "2^2" and p.fftParams.useBitReverse
    assert p.plfgAddress.base == 0x30000000 and p.plfgRAM.base == 0x30001000 and p.cfarAddress.base == 0x30002000
    assert p.fftAddress.base == 0x30000100 and p.magAddress.base == 0x30000200 and p.beatBytes == 4
    with pytest.raises(AssertionError):
        R.FFTParams.fixed(numPoints=1000)
    with pytest.raises(AssertionError):
        R.FFTParams.fixed(numPoints=8, keepMSBorLSB=[True])
    c = R.CFARParams(protoIn=R.FixedPoint(16, 0), protoThreshold=R.FixedPoint(16, 0), protoScaler=R.FixedPoint(16, 0),
                     leadLaggWindowSize=32, guardWindowSize=4, fftSize=512).to_c()
    assert (c.inWidth, c.leadLaggWindowSize, c.fftSize, c.CFARAlgorithm) == (16, 32, 512, 2)   # default CFARAlgorithm = GOSCACFARType (CFARParams.scala:29)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "rsp_chains_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                assert "rsp_oracle" not in text and "from oracle" not in text and "import oracle" not in text, f


def test_shard_rows_partitions_units():
    import rsp_chains_b200 as R
    for n_units in (1, 7, 8, 256, 1000):
        for world in (1, 2, 4, 8):
            spans = [R.shard_rows(n_units, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n_units
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["RSP_ROOT"])
import torch, torch.distributed as dist
import rsp_chains_b200 as R
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n_frames = 5
lo, hi = R.shard_rows(n_frames, world, rank)
# each rank "detects" one target per frame it owns: (chirp = frame, bin = 100 + frame, thr, cut)
det = torch.zeros((16, 4), dtype=torch.int32)
for k, f in enumerate(range(lo, hi)):
    det[k] = torch.tensor([f, 100 + f, 7 * f, 9 * f], dtype=torch.int32)
cnt = torch.tensor([hi - lo], dtype=torch.int32)
g = R.gather_detections(det, cnt)
assert g.shape == (n_frames, 5), g.shape
assert sorted(g[:, 1].tolist()) == list(range(n_frames))
for row in g.tolist():
    r, f, b, t, c = row
    assert (b, t, c) == (100 + f, 7 * f, 9 * f)
    assert R.shard_rows(n_frames, world, r)[0] <= f < R.shard_rows(n_frames, world, r)[1]
dist.barrier()
dist.destroy_process_group()
print("ok", rank)
"""


def test_gather_detections_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, RSP_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29631", str(script)], capture_output=True, text=True, env=env, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_bench_reference_arm_runs_on_cpu():
    """bench.py --impl reference prints one JSON line with the contract's keys (tiny run)"""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Msamples/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
