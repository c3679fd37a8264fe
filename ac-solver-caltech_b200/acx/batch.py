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
    """Device-resident state of N lockstep ACEnv instances (ac_env.py:53-89 fields, as arrays):
    state int8 [N, 2m] and one 32-bit word per env, meta = count_steps | len0 << 16 | len1 << 24.

    want_clip / want_status: allocate the optional per-row outputs (clipped float reward, row status); without
    them the step writes the 8-byte record only and errors are reported through `err_count`.
    track_actions: keep the action ring (uint8 [horizon, N]) and the finished-episode side buffer on the device --
    what `ACEnv.actions` / gymnasium's final_info hold in the reference (ac_env.py:92,110)."""

    def __init__(self, state, horizon, pool=None, want_clip=True, want_status=True, track_actions=False, fin_cap=None):
        n, row = state.shape
        self.n, self.m = n, row // 2
        self.horizon = int(horizon)
        if not 1 <= self.horizon <= 65535:
            raise ValueError("horizon_length must be in 1..65535 (count_steps is a 16-bit field of the env word)")
        dev = state.device
        self.state = state.contiguous()
        self.meta = torch.zeros(n, dtype=torch.int32, device=dev)
        # one 8-byte record per env written by the kernel: int32 reward | done | truncated | len0 | len1
        self.step_out = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
        self.reward = self.step_out.view(torch.int32)[:, 0]
        self.done = self.step_out[:, 4]
        self.truncated = self.step_out[:, 5]
        self.lens = self.step_out[:, 6:8]
        self.reward_clipped = torch.empty(n, dtype=torch.float32, device=dev) if want_clip else None
        self.status = torch.zeros(n, dtype=torch.uint8, device=dev) if want_status else None
        self.err_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.chain = torch.zeros(_lib.ENV_CHAIN_SLOTS, dtype=torch.int32, device=dev)
        self.pool = None if pool is None else pool.contiguous()
        self.t = 0  # number of steps taken (selects the ring row)
        self.ring = self.fin_count = self.fin_env = self.fin_steps = self.fin_state = None
        self.fin_cap = 0
        if track_actions:
            self.ring = torch.zeros((self.horizon, n), dtype=torch.uint8, device=dev)
            self.fin_cap = int(fin_cap if fin_cap is not None else n)
            self.fin_count = torch.zeros(1, dtype=torch.int32, device=dev)
            self.fin_env = torch.zeros(self.fin_cap, dtype=torch.int32, device=dev)
            self.fin_steps = torch.zeros(self.fin_cap, dtype=torch.int32, device=dev)
            self.fin_state = torch.zeros((self.fin_cap, row), dtype=torch.int8, device=dev)
        self._recertify()

    def _recertify(self, steps=None):
        """lengths -> meta (count_steps kept unless given) and certification of the states and the pool: when every
        row is certified the kernel skips its input checks"""
        keep = (self.meta & 0xFFFF) if steps is None else steps.to(torch.int32)
        self.certified = certify(self.state, self.step_out[:, 6:8])
        if self.certified and self.pool is not None:
            self.certified = certify(self.pool, None)
        stream = _prep(self.state.device)
        keep = keep.contiguous()
        _lib.check(_lib.lib().acx_env_meta_init(_ptr(self.state), _ptr(keep), _ptr(self.meta), self.n, self.m, stream),
                   "acx_env_meta_init")

    @property
    def steps(self):
        """count_steps of every env (a copy: the counters live in the low 16 bits of `meta`)"""
        return (self.meta & 0xFFFF).to(torch.int32)

    def set_steps(self, steps):
        self.meta.copy_((self.meta & ~0xFFFF) | (steps.to(torch.int32) & 0xFFFF))

    def set_state(self, state, steps=None):
        """Overwrite the env states (and optionally step counters); lengths and certification are redone."""
        self.state.copy_(state)
        self._recertify(steps)

    def set_row(self, i, row, steps=0):
        """env i <- row (ACEnv.reset with a starting_state, ac_env.py:113-129)"""
        self.state[i].copy_(row)
        m = self.m
        l0, l1 = int((row[:m] != 0).sum()), int((row[m:] != 0).sum())
        word = (int(steps) & 0xFFFF) | (l0 << 16) | (l1 << 24)
        self.meta[i] = word - (1 << 32) if word >= (1 << 31) else word
        self.step_out[i, 6] = l0
        self.step_out[i, 7] = l1
        if self.certified and not certify(row[None, :].contiguous()):
            self.certified = False  # an arbitrary starting state: back to the checked kernel


def env_step(buf, actions, next_idx=None, clip=(-10.0, 1000.0), chained=False, state_in=None, state_out=None, rec_out=None):
    """ACEnv.step over all envs in place (ac_env.py:91-111) + reset of finished envs from buf.pool.

    chained=True promises that `actions` (and next_idx) were complete before the previous env_step on this buffer
    was launched and that nothing else wrote its state in between: consecutive steps then overlap head-to-tail
    (acx.h: ACX_ENV_CHAINED).

    state_in / state_out / rec_out (int8 [N,2m], int8 [N,2m], uint8 [N,8]) redirect the rows read, the rows written and
    the step records, e.g. to rows t, t+1 of a rollout buffer: the step then needs no observation copy.  The caller
    keeps track of where the current states are (buf.state is not rebound)."""
    _chk(actions, torch.uint8, (buf.n,), "actions")
    for t, shape, name in ((state_in, buf.state.shape, "state_in"), (state_out, buf.state.shape, "state_out"),
                           (rec_out, buf.step_out.shape, "rec_out")):
        if t is not None:
            _chk(t, buf.state.dtype if name != "rec_out" else torch.uint8, shape, name)
    stream = _prep(buf.state.device)
    pool = buf.pool if next_idx is not None else None
    if next_idx is not None:
        _chk(next_idx, torch.int32, (buf.n,), "next_idx")
        if pool is None:
            raise ValueError("next_idx given but the env buffers hold no pool")
    io = _lib.EnvIO()
    io.state = _ptr(buf.state if state_in is None else state_in)
    io.state_out = _ptr(state_out)
    io.meta, io.actions = _ptr(buf.meta), _ptr(actions)
    io.step_out = _ptr(buf.step_out if rec_out is None else rec_out)
    io.reward_clipped, io.row_status, io.err_count = _ptr(buf.reward_clipped), _ptr(buf.status), _ptr(buf.err_count)
    io.pool, io.next_idx = _ptr(pool), _ptr(next_idx)
    io.action_ring, io.t, io.ring_len = _ptr(buf.ring), buf.t, buf.horizon
    io.fin_cap, io.fin_count, io.fin_env = buf.fin_cap, _ptr(buf.fin_count), _ptr(buf.fin_env)
    io.fin_steps, io.fin_state = _ptr(buf.fin_steps), _ptr(buf.fin_state)
    io.chain = _ptr(buf.chain)
    io.n, io.m, io.horizon = buf.n, buf.m, buf.horizon
    io.clip_lo, io.clip_hi = float(clip[0]), float(clip[1])
    io.flags = (_lib.ENV_CERTIFIED if buf.certified else 0) | (_lib.ENV_CHAINED if chained else 0)
    import ctypes

    rc = _lib.lib().acx_env_step(ctypes.addressof(io), stream)
    _lib.check(rc, "acx_env_step")
    buf.t += 1
    return buf
