"""tbf_multi (include/tbf_multi.h): the C++ multi-device host entry of SURVEY 8(e).

not-gpu: tbf_multi_plan (static block partition of the organ list sorted by table key) against
the Python plan the torchrun path uses (tunebfree_b200/sharding.py).
gpu: a plain C host (tests/c/test_multi.c, compiled here with gcc against libtbf_gpu.so) drives
two entries -- two GPUs when the box has them, two contexts on GPU 0 otherwise -- and must match
one context bit for bit."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from tunebfree_b200 import engine, sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plan_is_the_sorted_block_partition():
    L = engine.lib()
    L.tbf_multi_plan.argtypes = [C.c_int, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32),
                                 C.POINTER(C.c_uint32)]
    rng = np.random.RandomState(3)
    for n_dev, n in ((2, 24), (8, 16384), (4, 1001), (3, 5)):
        keys = rng.randint(0, 37, size=n).astype(np.uint64)
        dev = np.zeros(n, dtype=np.uint32)
        order = np.zeros(n, dtype=np.uint32)
        rc = L.tbf_multi_plan(n_dev, n, keys.ctypes.data_as(C.POINTER(C.c_uint64)),
                              dev.ctypes.data_as(C.POINTER(C.c_uint32)),
                              order.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert rc == 0
        assert sorted(order.tolist()) == list(range(n))
        assert np.all(np.diff(keys[order].astype(np.int64)) >= 0)          # sorted by key
        assert np.all(np.diff(dev[order].astype(np.int64)) >= 0)           # contiguous blocks
        per = -(-n // n_dev)
        assert np.bincount(dev, minlength=n_dev).max() <= per
        if n % n_dev == 0:
            # the same cut sharding.shard_plan_grouped makes for the one-process-per-GPU path
            for r in range(n_dev):
                mine, _ = sharding.shard_plan_grouped(r, n_dev, per, lambda g: int(keys[g]))
                assert sorted(mine.tolist()) == sorted(np.flatnonzero(dev == r).tolist())
    assert L.tbf_multi_plan(0, 4, None, None, None) != 0


def test_multi_symbols_exported():
    L = engine.lib()
    for name in ("tbf_multi_create", "tbf_multi_destroy", "tbf_multi_device_count", "tbf_multi_context",
                 "tbf_multi_last_error", "tbf_multi_plan", "tbf_multi_render_batch"):
        assert hasattr(L, name), name


@pytest.mark.gpu
def test_c_host_drives_two_devices(tmp_path):
    import torch
    exe = str(tmp_path / "test_multi")
    lib_dir = os.path.join(ROOT, "tunebfree_b200")
    subprocess.run(["gcc", "-O1", "-std=c11", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c", "test_multi.c"), "-o", exe,
                    "-L" + lib_dir, "-ltbf_gpu", "-Wl,-rpath," + lib_dir, "-lm"], check=True)
    second = 1 if torch.cuda.device_count() > 1 else 0
    r = subprocess.run([exe, "0", str(second)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "identical to one context" in r.stdout
