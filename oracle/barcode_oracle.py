"""ctypes front-end of oracle/barcode_oracle.c (restatement of the reference's C++ barcode programs)
plus helpers that run the compiled reference itself (oracle/_ref/, built by `make -C oracle ref`
where /root/reference exists).

TEST INFRASTRUCTURE ONLY -- same rules as ac_oracle.py: tests/, bench CPU-baseline legs and
__graft_entry__.smoke() may import it as the checker; the product package must not.
"""
import ctypes
import os
import subprocess
import tempfile

import numpy as np

import ac_oracle

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
_i8p = ctypes.POINTER(ctypes.c_int8)


class _Graph(ctypes.Structure):
    _fields_ = [("n_nodes", ctypes.c_int64), ("n_edges", ctypes.c_int64),
                ("node_size", ctypes.POINTER(ctypes.c_uint16)), ("edge", ctypes.POINTER(ctypes.c_int64)),
                ("edge_filt", ctypes.POINTER(ctypes.c_uint16)), ("rows", _i8p), ("row_m", ctypes.c_int)]


_ready = False


def _lib():
    global _ready
    ac_oracle.build()
    L = ac_oracle.lib()
    if not _ready:
        L.bco_move.argtypes = [_i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _i8p]
        L.bco_graph_bfs.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(_Graph)]
        L.bco_graph_free.argtypes = [ctypes.POINTER(_Graph)]
        L.bco_graph_free.restype = None
        L.bco_neighbourhood.argtypes = [_i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.bco_neighbourhood.restype = ctypes.c_int64
        _ready = True
    return L


def move(row, t, classic):
    """Presentation::move (AC_UTILS_as_sets.h:298-363) on a zero-padded row; returns (child row, size) or None."""
    row = np.ascontiguousarray(row, dtype=np.int8)
    out = np.zeros_like(row)
    sz = _lib().bco_move(row.ctypes.data_as(_i8p), len(row) // 2, int(t), int(bool(classic)), out.ctypes.data_as(_i8p))
    return (out, sz) if sz >= 0 else None


def graph_bfs(n_max, classic):
    """ac_bfs.cpp main(): dict(sizes u16[N], edges i64[E,2], efilt u16[E], rows i8[N, 2*(n_max-1)])."""
    g = _Graph()
    rc = _lib().bco_graph_bfs(int(n_max), int(bool(classic)), ctypes.byref(g))
    if rc != 0:
        raise RuntimeError("bco_graph_bfs failed: %d" % rc)
    try:
        n, e, m = g.n_nodes, g.n_edges, g.row_m
        return dict(
            sizes=np.ctypeslib.as_array(g.node_size, (n,)).copy(),
            edges=np.ctypeslib.as_array(g.edge, (max(e, 1), 2))[:e].copy(),
            efilt=np.ctypeslib.as_array(g.edge_filt, (max(e, 1),))[:e].copy(),
            rows=np.ctypeslib.as_array(g.rows, (n, 2 * m)).copy(),
        )
    finally:
        _lib().bco_graph_free(ctypes.byref(g))


def neighbourhood(row, radius, classic):
    """neibourhood() (neibourhoods.cpp:18-54)."""
    row = np.ascontiguousarray(row, dtype=np.int8)
    r = _lib().bco_neighbourhood(row.ctypes.data_as(_i8p), len(row) // 2, int(radius), int(bool(classic)))
    if r < 0:
        raise RuntimeError("bco_neighbourhood failed")
    return int(r)


# ---- the compiled reference (only where oracle/_ref exists) ----
def have_ref():
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in ("ac_bfs_classic", "ac_bfs_prime", "ref_neighbourhoods"))


def ref_ac_bfs_files(n_max, classic):
    """Run the reference ac_bfs and return its four output files as bytes."""
    exe = os.path.join(REF_DIR, "ac_bfs_classic" if classic else "ac_bfs_prime")
    with tempfile.TemporaryDirectory() as d:
        subprocess.check_call([exe, str(n_max)], cwd=d, stdout=subprocess.DEVNULL)
        return {k: open(os.path.join(d, "%s_%d" % (k, n_max)), "rb").read()
                for k in ("zero_simplices", "zero_filtrations", "one_simplices", "one_filtrations")}


def ref_neighbourhoods(rows, radius, classic):
    """Run the reference read_do_and_write on rows (lists of ints); returns the sizes."""
    exe = os.path.join(REF_DIR, "ref_neighbourhoods")
    with tempfile.TemporaryDirectory() as d:
        src, out = os.path.join(d, "in.txt"), os.path.join(d, "out.txt")
        with open(src, "w") as f:
            for r in rows:
                f.write("[" + ", ".join(str(int(v)) for v in r) + "]\n")
        subprocess.check_call([exe, src, out, str(radius), "1" if classic else "0"], stdout=subprocess.DEVNULL)
        return [int(x) for x in open(out).read().split()]
