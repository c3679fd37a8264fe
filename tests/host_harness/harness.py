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
    deps = glob.glob(os.path.join(ROOT, "ac-solver-caltech_b200", "csrc", "*.cuh")) + [os.path.join(HERE, "acx_host.cu"),
                                                                                      os.path.join(ROOT, "include", "acx.h")]
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
        _P, _I64, _I32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
        L.acxh_graph_set_order.argtypes = [_I32]
        L.acxh_graph_set_order.restype = None
        L.acxh_graph_scan_bytes.argtypes = [_I64]
        L.acxh_graph_scan_bytes.restype = _I64
        L.acxh_graph_step.argtypes = [_P, _P]
        L.acxh_graph_pack.argtypes = [_P, _I64, _I32, _P, _I32, _P, _P, _P]
        L.acxh_graph_unpack.argtypes = [_P, _I64, _I32, _I32, _P, _P, _P]
        L.acxh_graph_rehash.argtypes = [_P, _I64, _I32, _P, _I64, _P]
        L.acxh_graph_count_instances.argtypes = [_P, _I64, _I32, _P, _I64, _P]
        L.acxh_ms_enumerate.argtypes = [_I32, _I64, _I64, _P, _P, _P]
        L.acxh_graph_move.argtypes = [_P, _I64, _I32, _P, _I32, _I32, _I32, _P, _P]
        L.acxh_path_expand_cyclic.argtypes = [_P, _P, _P, _I64, _I32, _I32, _P, _P, _I32, _P, _P, _P, _P]
        _lib = L
    return _lib


class HostGraphBackend:
    """Stand-in for acx.graph._CudaBackend in the CPU suite: the same csrc/acx_graph.cuh functors and
    graph_step<> template, run sequentially on CPU tensors (order: 0 forward, 1 reverse, 2 permuted)."""

    def __init__(self, order=0):
        import torch

        self.dev = torch.device("cpu")
        self.stream = None
        self.L = lib()
        self.order = order

    def call(self, name, *args):
        self.L.acxh_graph_set_order(self.order)
        rc = getattr(self.L, "acxh_graph_" + name)(*args, None)
        assert rc == 0, (name, rc)

    def scan_bytes(self, n):
        return int(self.L.acxh_graph_scan_bytes(n))


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


class HostMsBackend:
    """Stand-in for the CUDA enumeration step of ac_solver.search.miller_schupp (csrc/acx_ms.cuh on the CPU)."""

    def __init__(self):
        import torch

        self.dev = torch.device("cpu")
        self.L = lib()

    def enumerate(self, w_len, first, count, key, rel):
        assert self.L.acxh_ms_enumerate(w_len, first, count, key.data_ptr(), rel.data_ptr(), None) == 0


class HostExpandBackend:
    """Stand-in for acx.solution_cache._CudaExpand (csrc/acx_cyclic_expand.cuh on the CPU)."""

    def __init__(self):
        import torch

        self.dev = torch.device("cpu")
        self.L = lib()

    def expand(self, *args):
        assert self.L.acxh_path_expand_cyclic(*args, None) == 0


def graph_move(rows, moves, classic, width_words, allow_empty=False):
    """g_move + sort of csrc/acx_graph.cuh on byte rows [n, 2m]: (children int8 [n, 2m], status int32 [n])
    (status 0 ok, 1 relator over the plane capacity, 2 empty relator, -1 row outside the domain)."""
    rows = np.ascontiguousarray(rows, dtype=np.int8)
    mv = np.ascontiguousarray(moves, dtype=np.uint8)
    out = np.zeros_like(rows)
    st = np.zeros(len(rows), dtype=np.int32)
    rc = lib().acxh_graph_move(rows.ctypes.data, len(rows), rows.shape[1] // 2, mv.ctypes.data, int(bool(classic)),
                               int(bool(allow_empty)), width_words, out.ctypes.data, st.ctypes.data)
    assert rc == 0
    return out, st
