"""Builds and loads tests/host_harness/libacx_host.so: the product's __host__ __device__ row
arithmetic (csrc/acx_rows.cuh) compiled for the CPU.  TEST INFRASTRUCTURE -- lets the CPU suite
check the plane algebra against the oracle; the product never loads it."""
import ctypes
import glob
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "libacx_host.so")
_i8p, _u8p = ctypes.POINTER(ctypes.c_int8), ctypes.POINTER(ctypes.c_uint8)
_lib = None


def build():
    deps = glob.glob(os.path.join(ROOT, "ac-solver-caltech_b200", "csrc", "*.cuh")) + [os.path.join(HERE, "acx_host.cu")]
    if os.path.exists(SO) and os.path.getmtime(SO) >= max(os.path.getmtime(d) for d in deps):
        return SO
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2",
                           "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-o", SO, os.path.join(HERE, "acx_host.cu")])
    return SO


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        L.acxh_move_batch.argtypes = [_i8p, _u8p, _i8p, _u8p, _u8p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_int]
        _lib = L
    return _lib


def move_batch(states, moves, cyclical, variant=0, inplace=1):
    s = np.ascontiguousarray(states, dtype=np.int8)
    mv = np.ascontiguousarray(moves, dtype=np.uint8)
    n, row = s.shape
    out = np.empty_like(s)
    lens = np.empty((n, 2), np.uint8)
    st = np.empty(n, np.uint8)
    rc = lib().acxh_move_batch(s.ctypes.data_as(_i8p), mv.ctypes.data_as(_u8p), out.ctypes.data_as(_i8p),
                               lens.ctypes.data_as(_u8p), st.ctypes.data_as(_u8p), n, row // 2, int(cyclical),
                               variant, inplace)
    assert rc == 0
    return out, lens, st
