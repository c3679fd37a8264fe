"""ctypes loader of libacx.so (built in-tree by ../build.py).

There is no fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libacx.so")
_lib = None

c_i8p = ctypes.c_void_p  # device pointers are passed as integers
_I64, _I32, _F32, _PTR = ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_void_p

_SIGNATURES = {
    "acx_last_error": (ctypes.c_char_p, []),
    "acx_version": (_I32, []),
    "acx_device_count": (_I32, []),
    "acx_set_device": (_I32, [_I32]),
    "acx_move_batch": (_I32, [_PTR, _PTR, _PTR, _PTR, _PTR, _I64, _I32, _I32, _PTR]),
    "acx_expand12": (_I32, [_PTR, _PTR, _PTR, _PTR, _I64, _I32, _I32, _PTR]),
    "acx_search_workspace_bytes": (_I64, [_I32, _I32, _I64]),
    "acx_search_batch": (_I32, [_PTR, _I32, _I32, _I64, _I32, _I32, _PTR, _I64, _PTR, _PTR, _I32, _PTR]),
    "acx_rows_primitive": (_I32, [_PTR, _PTR, _PTR, _PTR, _PTR, _I64, _I32, _I32, _PTR]),
    "acx_bfs_seed": (_I32, [_PTR, _I32, _PTR, _I64, _I64, _PTR]),
    "acx_bfs_probe": (_I32, [_PTR, _PTR, _PTR, _PTR, _I64, _I32, _PTR, _I64, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_bfs_commit": (_I32, [_PTR, _PTR, _PTR, _PTR, _PTR, _PTR, _I64, _I64, _I64, _I32, _PTR, _PTR, _PTR, _PTR, _PTR,
                              _I64, _PTR]),
    "acx_rows_hash": (_I32, [_PTR, _I64, _I32, _PTR, _PTR]),
    "acx_bfs_record_bytes": (_I32, [_I32]),
    "acx_bfs_expand_bucket": (_I32, [_PTR, _PTR, _PTR, _I64, _I64, _I32, _I32, _I32, _I64, _PTR, _PTR, _PTR]),
    "acx_bfs_bucket": (_I32, [_PTR, _PTR, _PTR, _PTR, _I64, _I32, _I32, _PTR, _PTR, _PTR, _PTR, _PTR, _I32, _PTR]),
    "acx_bfs_probe_packed": (_I32, [_PTR, _PTR, _I32, _I64, _I32, _PTR, _I64, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_bfs_commit_packed": (_I32, [_PTR, _PTR, _I32, _I64, _I32, _PTR, _PTR, _PTR, _I64, _I64, _PTR, _PTR, _PTR, _PTR, _PTR,
                                     _PTR, _PTR, _PTR]),
    "acx_pbfs_pack": (_I32, [_PTR, _I64, _I32, _I32, _PTR, _PTR, _PTR]),
    "acx_pbfs_unpack": (_I32, [_PTR, _I64, _I32, _PTR, _PTR]),
    "acx_pbfs_hash": (_I32, [_PTR, _I64, _PTR, _PTR]),
    "acx_pbfs_seed": (_I32, [_PTR, _PTR, _I64, _I64, _PTR]),
    "acx_pbfs_expand": (_I32, [_PTR, _PTR, _PTR, _I64, _I64, _I32, _I32, _I32, _I64, _PTR, _PTR, _I32, _PTR]),
    "acx_pbfs_probe": (_I32, [_PTR, _PTR, _I32, _I64, _PTR, _I64, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_pbfs_commit": (_I32, [_PTR, _PTR, _I32, _I64, _PTR, _PTR, _PTR, _I64, _I64, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR,
                               _PTR]),
    "acx_pbfs_expand2": (_I32, [_PTR, _PTR, _PTR, _PTR, _I64, _I64, _I32, _I32, _I32, _I64, _PTR, _PTR, _PTR, _PTR]),
    "acx_pbfs_probe3": (_I32, [_PTR, _PTR, _I64, _I32, _I64, _PTR, _I64, _PTR, _I64, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_pbfs_decide": (_I64, [_PTR, _PTR, _I64, _PTR, _I64, _I64, _PTR, _PTR, _I32, _I32, _PTR, _I64, _PTR]),
    "acx_pbfs_commit3": (_I32, [_PTR, _PTR, _I64, _I64, _PTR, _PTR, _I64, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_bfs_mask_scan": (_I64, [_PTR, _I64, _PTR, _PTR, _PTR, _I64, _PTR]),
    "acx_bfs_mask_scan_static": (_I32, [_PTR, _I64, _PTR, _PTR, _PTR]),
    "acx_bfs_window_partition": (_I32, [_PTR, _I64, _I64, _I64, _I64, _PTR, _PTR, _PTR, _I64, _PTR, _PTR, _I32, _PTR]),
    "acx_path_replay": (_I32, [_PTR, _PTR, _PTR, _I64, _I32, _I32, _I32, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR, _PTR]),
    "acx_rows_features": (_I32, [_PTR, _I64, _I32, _PTR, _PTR]),
    "acx_path_expand_cyclic": (_I32, [_PTR, _PTR, _PTR, _I64, _I32, _I32, _PTR, _PTR, _I32, _PTR, _PTR, _PTR, _PTR]),
    "acx_peer_alloc": (_I32, [_I64, _PTR]),
    "acx_peer_free": (_I32, [_PTR]),
    "acx_peer_export": (_I32, [_PTR, _PTR]),
    "acx_peer_open": (_I32, [_PTR, _PTR]),
    "acx_peer_close": (_I32, [_PTR]),
    "acx_env_step": (_I32, [_PTR, _PTR]),
    "acx_env_meta_init": (_I32, [_PTR, _PTR, _PTR, _I64, _I32, _PTR]),
    "acx_rows_certify": (_I32, [_PTR, _I64, _I32, _PTR, _PTR, _PTR]),
    "acx_graph_scan_bytes": (_I64, [_I64]),
    "acx_graph_step": (_I32, [_PTR, _PTR]),
    "acx_graph_pack": (_I32, [_PTR, _I64, _I32, _PTR, _I32, _PTR, _PTR, _PTR]),
    "acx_graph_unpack": (_I32, [_PTR, _I64, _I32, _I32, _PTR, _PTR, _PTR]),
    "acx_graph_rehash": (_I32, [_PTR, _I64, _I32, _PTR, _I64, _PTR]),
    "acx_graph_count_instances": (_I32, [_PTR, _I64, _I32, _PTR, _I64, _PTR]),
    "acx_ms_enumerate": (_I32, [_I32, _I64, _I64, _PTR, _PTR, _PTR]),
}


class EnvIO(ctypes.Structure):
    """acx_env_io of include/acx.h (field for field)"""

    _fields_ = [
        ("state", _PTR), ("state_out", _PTR), ("meta", _PTR), ("actions", _PTR), ("step_out", _PTR), ("reward_clipped", _PTR),
        ("row_status", _PTR), ("err_count", _PTR), ("pool", _PTR), ("next_idx", _PTR), ("action_ring", _PTR),
        ("t", _I64), ("ring_len", ctypes.c_int32), ("fin_cap", ctypes.c_int32), ("fin_count", _PTR), ("fin_env", _PTR),
        ("fin_steps", _PTR), ("fin_state", _PTR), ("chain", _PTR), ("n", _I64), ("m", ctypes.c_int32),
        ("horizon", ctypes.c_int32), ("clip_lo", _F32), ("clip_hi", _F32), ("flags", ctypes.c_int32),
    ]


ENV_CERTIFIED, ENV_CHAINED, ENV_CHAIN_SLOTS = 1, 2, 4096
MASK_SCAN_STATIC_MAX_N, MASK_SCAN_STATIC_BYTES = 1 << 22, 256 * 1024  # include/acx.h


class AcxError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python ac-solver-caltech_b200/build.py` "
                "(there is no CPU or PyTorch fallback for the AC-move engine)")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def exported_symbols():
    return list(_SIGNATURES)


def check(rc, what):
    if rc != 0:
        msg = lib().acx_last_error()
        raise AcxError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
