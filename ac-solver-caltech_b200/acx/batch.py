"""Batch entry points on torch CUDA tensors (thin wrappers over the C ABI in include/acx.h).

Tensors are used for device memory and streams only; all arithmetic happens in libacx.so.
"""
import torch

from . import _lib

_current_dev = [None]


def _prep(device):
    if device.type != "cuda":
        raise _lib.AcxError("acx needs CUDA tensors: there is no CPU implementation of the AC-move engine")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if _current_dev[0] != idx:
        _lib.check(_lib.lib().acx_set_device(idx), "acx_set_device")
        _current_dev[0] = idx
    return torch.cuda.current_stream(idx).cuda_stream


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _chk(t, dtype, shape, name):
    if t.dtype != dtype or not t.is_contiguous() or tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name}: expected contiguous {dtype} of shape {tuple(shape)}, got {t.dtype} {tuple(t.shape)}")


def move_batch(states, moves, cyclical=True, out=None, want_lens=True, want_status=True):
    """ACMove over rows: states int8 [N,2m], moves uint8 [N] -> (next, lens u8 [N,2], status u8 [N])."""
    n, row = states.shape
    m = row // 2
    _chk(states, torch.int8, (n, 2 * m), "states")
    _chk(moves, torch.uint8, (n,), "moves")
    stream = _prep(states.device)
    if out is None:
        out = torch.empty_like(states)
    else:
        _chk(out, torch.int8, (n, 2 * m), "out")
    lens = torch.empty((n, 2), dtype=torch.uint8, device=states.device) if want_lens else None
    status = torch.empty((n,), dtype=torch.uint8, device=states.device) if want_status else None
    rc = _lib.lib().acx_move_batch(_ptr(states), _ptr(moves), _ptr(out), _ptr(lens), _ptr(status), n, m,
                                   int(bool(cyclical)), stream)
    _lib.check(rc, "acx_move_batch")
    return out, lens, status


def expand12(parents, cyclical=False, want_lens=True, want_status=True):
    """All 12 successors: parents int8 [F,2m] -> (children [F,12,2m], lens u8 [F,12,2], status u8 [F,12])."""
    f, row = parents.shape
    m = row // 2
    _chk(parents, torch.int8, (f, 2 * m), "parents")
    stream = _prep(parents.device)
    children = torch.empty((f, 12, row), dtype=torch.int8, device=parents.device)
    lens = torch.empty((f, 12, 2), dtype=torch.uint8, device=parents.device) if want_lens else None
    status = torch.empty((f, 12), dtype=torch.uint8, device=parents.device) if want_status else None
    rc = _lib.lib().acx_expand12(_ptr(parents), _ptr(children), _ptr(lens), _ptr(status), f, m,
                                 int(bool(cyclical)), stream)
    _lib.check(rc, "acx_expand12")
    return children, lens, status


def certify(rows, lens_out=None):
    """acx_rows_certify: True iff every row is certified; writes the lengths into lens_out (u8 [N,2] view)."""
    n, row = rows.shape
    _chk(rows, torch.int8, (n, row), "rows")
    stream = _prep(rows.device)
    lens = torch.empty((n, 2), dtype=torch.uint8, device=rows.device)
    cert = torch.empty((n,), dtype=torch.uint8, device=rows.device)
    rc = _lib.lib().acx_rows_certify(_ptr(rows), n, row // 2, _ptr(lens), _ptr(cert), stream)
    _lib.check(rc, "acx_rows_certify")
    if lens_out is not None:
        lens_out.copy_(lens)
    return bool(cert.all().item()) if n else True


class EnvBuffers:
    """Device-resident state of N lockstep ACEnv instances (ac_env.py:53-89 fields, as arrays)."""

    def __init__(self, state, horizon, pool=None):
        n, row = state.shape
        self.n, self.m = n, row // 2
        self.horizon = int(horizon)
        dev = state.device
        self.state = state.contiguous()
        self.steps = torch.zeros(n, dtype=torch.int32, device=dev)
        # one 8-byte record per env written by the kernel: int32 reward | done | truncated | len0 | len1
        self.step_out = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
        self.reward = self.step_out.view(torch.int32)[:, 0]
        self.done = self.step_out[:, 4]
        self.truncated = self.step_out[:, 5]
        self.lens = self.step_out[:, 6:8]
        self.reward_clipped = torch.empty(n, dtype=torch.float32, device=dev)
        # certification: when every state and pool row is certified the kernel skips its input checks
        self.certified = certify(self.state, self.step_out[:, 6:8])
        if self.certified and pool is not None:
            self.certified = certify(pool.contiguous(), None)
        self.status = torch.empty(n, dtype=torch.uint8, device=dev)
        self.pool = None if pool is None else pool.contiguous()


def _set_state(self, state, steps=None):
    """Overwrite the env states (and optionally step counters); lengths and certification are redone."""
    self.state.copy_(state)
    if steps is not None:
        self.steps.copy_(steps)
    self.certified = certify(self.state, self.step_out[:, 6:8])
    if self.certified and self.pool is not None:
        self.certified = certify(self.pool, None)


EnvBuffers.set_state = _set_state


def env_step(buf, actions, next_idx=None, clip=(-10.0, 1000.0)):
    """ACEnv.step over all envs in place (ac_env.py:91-111) + reset of finished envs from buf.pool."""
    _chk(actions, torch.uint8, (buf.n,), "actions")
    stream = _prep(buf.state.device)
    pool = buf.pool if next_idx is not None else None
    if next_idx is not None:
        _chk(next_idx, torch.int32, (buf.n,), "next_idx")
        if pool is None:
            raise ValueError("next_idx given but the env buffers hold no pool")
    rc = _lib.lib().acx_env_step(_ptr(buf.state), _ptr(buf.steps), _ptr(actions), _ptr(pool), _ptr(next_idx),
                                 _ptr(buf.step_out), _ptr(buf.reward_clipped), _ptr(buf.status), buf.n, buf.m,
                                 buf.horizon, float(clip[0]), float(clip[1]),
                                 1 if buf.certified else 0, stream)
    _lib.check(rc, "acx_env_step")
    return buf
