"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/vxrt.h declares,
validates arguments, and fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "vxrt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vx_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(vx):
    L = vx.lib()
    syms = header_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(L, s), "libvxrt.so does not export %s" % s
    assert sorted(vx.ABI_SYMBOLS) == syms
    assert L.vx_version() == 100


def test_argument_validation_needs_no_gpu(vx):
    L = vx.lib()
    h = C.c_void_p()
    # factor must be a multiple of 8, dims multiples of factor, coarse dims multiples of 8
    assert L.vx_world_build_noise(C.byref(h), 64, 64, 64, 4, 0, 0, 0, 0) == -1
    assert b"factor" in L.vx_last_error()
    assert L.vx_world_build_noise(C.byref(h), 60, 64, 64, 8, 0, 0, 0, 0) == -1
    assert L.vx_world_build_noise(C.byref(h), 32, 64, 64, 8, 0, 0, 0, 0) == -1
    assert L.vx_world_from_dense_bits(C.byref(h), None, 64, 64, 64, 8, 0) == -1


def test_get_directions_matches_oracle(vx, oracle):
    for e in [(0, 0, 0), (-0.35, 0, 0), (-0.05, 0.8, 0), (0.3, -2.0, 0.1)]:
        a = vx.get_directions(e)
        b = oracle.get_directions(e)
        for p, q in zip(a, b):
            assert np.array_equal(p, q)
    f, u, r = vx.get_directions((0, 0, 0))
    assert np.allclose(f, [0, 0, -1]) and np.allclose(u, [0, -1, 0]) and np.allclose(r, [1, 0, 0])


def test_postprocess_matches_reference_semantics(vx):
    o = np.array([[0, 0, 0], [1, 2, 3]], np.float32)
    p = np.array([[3, 4, 0], [np.inf, np.inf, np.inf]], np.float32)
    valid = np.zeros(2, np.uint8)
    dist = np.zeros(2, np.float32)
    assert vx.lib().vx_hits_postprocess(o.ctypes.data, p.ctypes.data, 2, valid.ctypes.data, dist.ctypes.data) == 0
    assert valid.tolist() == [1, 0] and dist[0] == 5.0 and np.isinf(dist[1])


def test_no_cpu_fallback_without_device(vx):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(vx.VxError):
        vx.World.from_noise(64, 64, 64, 8)


def test_bench_stdout_is_one_json_line_whatever_libraries_print():
    """bench.py's contract is ONE JSON line on stdout; NCCL writes its version banner to file descriptor 1 from every rank
    (at NCCL_DEBUG=WARN too).  bench.claim_stdout() points fd 1 at stderr and emit() writes to the original descriptor."""
    import json
    import subprocess
    import sys
    code = ("import os, sys; sys.argv = ['bench.py']; import bench; bench.claim_stdout(); os.write(1, b'NCCL version 2.28.9+cuda12.9\\n'); "
            "print('a library print'); bench.emit({'metric': 'Mrays/s', 'value': 1.0})")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    assert len(lines) == 1 and json.loads(lines[0])["value"] == 1.0
    assert "NCCL version" in r.stderr and "a library print" in r.stderr
