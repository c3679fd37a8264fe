"""ctypes front-end of the CPU oracle (oracle/ac_oracle.c, oracle/ac_search_oracle.c).

TEST INFRASTRUCTURE ONLY.  May be imported from tests/, from bench.py's
cpu_baseline / --impl reference legs and from __graft_entry__.smoke() -- as the
checker, never as the thing shipped.  The product package (acx, ac_solver under
ac-solver-caltech_b200/) must not import this module.

Function names follow the reference (ac_solver/envs/ac_moves.py, utils.py,
ac_env.py, search/greedy.py, search/breadth_first.py).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

_i8p = ctypes.POINTER(ctypes.c_int8)
_u8p = ctypes.POINTER(ctypes.c_uint8)
_i32p = ctypes.POINTER(ctypes.c_int32)
_i64p = ctypes.POINTER(ctypes.c_int64)


def build(force=False):
    """Compile liboracle.so with the Makefile next to this file."""
    srcs = [os.path.join(_HERE, f) for f in ("ac_oracle.c", "ac_search_oracle.c", "barcode_oracle.c", "ac_oracle.h")]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs
    )
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        L.orc_is_valid.argtypes = [_i8p, ctypes.c_int]
        L.orc_is_trivial.argtypes = [_i8p, ctypes.c_int]
        L.orc_simplify_relator.argtypes = [_i8p, ctypes.c_int, ctypes.c_int]
        L.orc_simplify_presentation.argtypes = [_i8p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.orc_concatenate.argtypes = [_i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_concatenate.restype = None
        L.orc_conjugate.argtypes = [_i8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_conjugate.restype = None
        L.orc_acmove.argtypes = [ctypes.c_int, _i8p, _i8p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.orc_acmove_batch.argtypes = [_i8p, _u8p, _i8p, _i32p, _u8p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_acmove_batch.restype = None
        L.orc_expand12_batch.argtypes = [_i8p, _i8p, _i32p, _u8p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_expand12_batch.restype = None
        L.orc_env_step_batch.argtypes = [_i8p, _i32p, _u8p, _i8p, _i32p, _i32p, _u8p, _u8p, _u8p, _u8p,
                                         ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_env_step_batch.restype = None
        for f in (L.orc_greedy, L.orc_bfs):
            f.argtypes = [_i8p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, _i32p, _i32p, ctypes.c_int,
                          ctypes.POINTER(ctypes.c_int), _i64p, _i64p]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def _row(presentation):
    a = np.ascontiguousarray(np.asarray(presentation), dtype=np.int8)
    assert a.ndim == 1 and a.size % 2 == 0
    return a


# ---- scalar API (reference signatures) --------------------------------------
def is_array_valid_presentation(array):
    a = np.asarray(array)
    if a.size % 2:
        return False
    a = _row(a)
    return bool(lib().orc_is_valid(_p(a, _i8p), a.size // 2))


def is_presentation_trivial(presentation):
    a = np.asarray(presentation)
    if a.size % 2:
        return False
    a = _row(a)
    return bool(lib().orc_is_trivial(_p(a, _i8p), a.size // 2))


def simplify_relator(relator, max_relator_length, cyclical=False, padded=True):
    """utils.py:178-243.  The relator may be shorter or longer than max_relator_length."""
    r = np.asarray(relator)
    width = max(len(r), max_relator_length)
    buf = np.zeros(width, dtype=np.int8)
    buf[: len(r)] = r
    n = lib().orc_simplify_relator(_p(buf, _i8p), width, int(bool(cyclical)))
    assert max_relator_length >= n, "Increase max length!"
    if padded:
        out = buf[:max_relator_length].copy()
    else:  # np.delete keeps whatever zero padding the input carried (utils.py:214, :230)
        out = buf[: len(r) - (int(np.count_nonzero(r)) - n)].copy()
    return out.astype(r.dtype if r.size else np.int8), n


def simplify_presentation(presentation, max_relator_length, lengths_of_words=None, cyclical=True):
    a = _row(presentation).copy()
    lens = (ctypes.c_int * 2)()
    st = lib().orc_simplify_presentation(_p(a, _i8p), max_relator_length, int(bool(cyclical)), lens)
    assert st == 0, f"{presentation} is not a valid presentation."
    return a.astype(np.asarray(presentation).dtype), [lens[0], lens[1]]


def concatenate_relators(presentation, max_relator_length, i, j, sign, lengths=None):
    assert i in (0, 1) and j == 1 - i and sign in (1, -1)
    a = _row(presentation).copy()
    lib().orc_concatenate(_p(a, _i8p), max_relator_length, i, sign)
    m = max_relator_length
    return a.astype(np.asarray(presentation).dtype), [int(np.count_nonzero(a[:m])), int(np.count_nonzero(a[m:]))]


def conjugate(presentation, max_relator_length, i, j, sign, lengths=None):
    assert i in (0, 1) and j in (1, 2) and sign in (1, -1)
    a = _row(presentation).copy()
    lib().orc_conjugate(_p(a, _i8p), max_relator_length, i, j, sign)
    m = max_relator_length
    return a.astype(np.asarray(presentation).dtype), [int(np.count_nonzero(a[:m])), int(np.count_nonzero(a[m:]))]


def ACMove(move_id, presentation, max_relator_length, lengths=None, cyclical=True):
    assert move_id in range(0, 12)
    a = _row(presentation)
    out = np.empty_like(a)
    lens = (ctypes.c_int * 2)()
    st = lib().orc_acmove(int(move_id), _p(a, _i8p), _p(out, _i8p), max_relator_length, int(bool(cyclical)), lens)
    assert st == 0, f"row is not a valid presentation (status {st})"
    return out.astype(np.asarray(presentation).dtype), [lens[0], lens[1]]


# ---- batches ----------------------------------------------------------------
def acmove_batch(states, moves, cyclical=True, threads=1):
    """states [N,2m] int8, moves [N] -> (next [N,2m] int8, lens [N,2] int32, status [N] uint8)."""
    s = np.ascontiguousarray(states, dtype=np.int8)
    mv = np.ascontiguousarray(moves, dtype=np.uint8)
    n, row = s.shape
    out = np.empty_like(s)
    lens = np.empty((n, 2), dtype=np.int32)
    status = np.empty(n, dtype=np.uint8)
    lib().orc_acmove_batch(_p(s, _i8p), _p(mv, _u8p), _p(out, _i8p), _p(lens, _i32p), _p(status, _u8p),
                           n, row // 2, int(bool(cyclical)), threads)
    return out, lens, status


def expand12_batch(states, cyclical=False, threads=1):
    """states [N,2m] -> (children [N,12,2m], lens [N,12,2], status [N,12])."""
    s = np.ascontiguousarray(states, dtype=np.int8)
    n, row = s.shape
    out = np.empty((n, 12, row), dtype=np.int8)
    lens = np.empty((n, 12, 2), dtype=np.int32)
    status = np.empty((n, 12), dtype=np.uint8)
    lib().orc_expand12_batch(_p(s, _i8p), _p(out, _i8p), _p(lens, _i32p), _p(status, _u8p),
                             n, row // 2, int(bool(cyclical)), threads)
    return out, lens, status


def env_step_batch(state, steps, actions, horizon, pool=None, next_idx=None, threads=1):
    """In place on state [N,2m] int8 and steps [N] int32 (ac_env.py:91-129).
    Returns (reward int32, done u8, truncated u8, lens u8 [N,2], status u8)."""
    assert state.dtype == np.int8 and state.flags.c_contiguous
    assert steps.dtype == np.int32 and steps.flags.c_contiguous
    n, row = state.shape
    act = np.ascontiguousarray(actions, dtype=np.uint8)
    reward = np.empty(n, dtype=np.int32)
    done = np.empty(n, dtype=np.uint8)
    trunc = np.empty(n, dtype=np.uint8)
    lens = np.empty((n, 2), dtype=np.uint8)
    status = np.empty(n, dtype=np.uint8)
    if pool is not None:
        pool = np.ascontiguousarray(pool, dtype=np.int8)
        next_idx = np.ascontiguousarray(next_idx, dtype=np.int32)
        pp, ip = _p(pool, _i8p), _p(next_idx, _i32p)
    else:
        pp, ip = None, None
    lib().orc_env_step_batch(_p(state, _i8p), _p(steps, _i32p), _p(act, _u8p), pp, ip,
                             _p(reward, _i32p), _p(done, _u8p), _p(trunc, _u8p), _p(lens, _u8p),
                             _p(status, _u8p), n, row // 2, int(horizon), threads)
    return reward, done, trunc, lens, status


# ---- searches ---------------------------------------------------------------
def _search(fn, presentation, max_nodes_to_explore, cyclical, stats):
    a = _row(presentation)
    cap = 1 << 16
    pa = np.empty(cap, dtype=np.int32)
    pl = np.empty(cap, dtype=np.int32)
    pn = ctypes.c_int(0)
    seen = ctypes.c_int64(0)
    expanded = ctypes.c_int64(0)
    rc = fn(_p(a, _i8p), a.size // 2, int(max_nodes_to_explore), int(bool(cyclical)),
            _p(pa, _i32p), _p(pl, _i32p), cap, ctypes.byref(pn), ctypes.byref(seen), ctypes.byref(expanded))
    if rc == -1:
        raise AssertionError("a relator became empty during the search")
    if rc < 0:
        raise RuntimeError(f"oracle search failed ({rc})")
    path = [(int(x), int(y)) for x, y in zip(pa[: pn.value], pl[: pn.value])]
    if stats is not None:
        stats["n_seen"] = seen.value
        stats["n_expanded"] = expanded.value
    return bool(rc), path


def greedy_search(presentation, max_nodes_to_explore=10000, verbose=False,
                  cyclically_reduce_after_moves=False, stats=None):
    return _search(lib().orc_greedy, presentation, max_nodes_to_explore, cyclically_reduce_after_moves, stats)


def bfs(presentation, max_nodes_to_explore=10000, verbose=False,
        cyclically_reduce_after_moves=False, stats=None):
    ok, path = _search(lib().orc_bfs, presentation, max_nodes_to_explore, cyclically_reduce_after_moves, stats)
    return ok, (path if ok else None)
