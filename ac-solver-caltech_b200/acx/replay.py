"""Batched path replay on the GPU: ACMove along a time axis, all intermediate states.

    out = replay_paths(roots, action_lists, cyclical=True)      # many paths in one launch
    state = replay_final_state(action_list, initial_state)      # agents/training.py:74-91, same signature
    path, detailed = actions_to_path(action_list, initial_state)  # agents/training.py:94-135, same return value

A step that would empty a relator raises AssertionError, as ACMove does (utils.py:264-266)."""
import numpy as np
import torch

from . import _lib
from .batch import _prep, _ptr


class Replay:
    """states int8 [P, T, 2m], lens uint8 [P, T, 2] (rows at t >= path_len[p] are undefined), final int8 [P, 2m],
    final_lens uint8 [P, 2], status uint8 [P], fail_step int32 [P], path_len int32 [P] -- CUDA tensors."""

    def __init__(self, **kw):
        self.__dict__.update(kw)


def replay_paths(roots, action_lists, cyclical=True, device="cuda:0", keep_states=True):
    roots = np.ascontiguousarray(np.asarray(roots), dtype=np.int8)
    if roots.ndim != 2 or roots.shape[1] % 2:
        raise ValueError("roots must be [P, 2*max_relator_length]")
    P, rowb = roots.shape
    if len(action_lists) != P:
        raise ValueError("one action list per root")
    T = max([len(a) for a in action_lists], default=0)
    acts = np.zeros((P, max(T, 1)), dtype=np.uint8)
    plen = np.zeros((P,), dtype=np.int32)
    for p, a in enumerate(action_lists):
        a = np.asarray(a, dtype=np.int64)
        if a.size and (a.min() < 0 or a.max() > 11):
            raise ValueError("move ids are 0..11")
        acts[p, : a.size] = a
        plen[p] = a.size
    dev = torch.device(device)
    stream = _prep(dev)
    d_roots, d_acts, d_len = (torch.from_numpy(x).to(dev) for x in (roots, acts, plen))
    Tm = acts.shape[1]
    states = torch.empty((P, Tm, rowb), dtype=torch.int8, device=dev) if keep_states else None
    lens = torch.empty((P, Tm, 2), dtype=torch.uint8, device=dev)
    final = torch.empty((P, rowb), dtype=torch.int8, device=dev)
    flen = torch.empty((P, 2), dtype=torch.uint8, device=dev)
    fail = torch.empty((P,), dtype=torch.int32, device=dev)
    status = torch.empty((P,), dtype=torch.uint8, device=dev)
    _lib.check(_lib.lib().acx_path_replay(_ptr(d_roots), _ptr(d_acts), _ptr(d_len), P, Tm if T else 0, rowb // 2, int(bool(cyclical)),
                                          _ptr(states) if keep_states else 0, _ptr(lens), _ptr(final), _ptr(flen), _ptr(fail),
                                          _ptr(status), stream), "acx_path_replay")
    return Replay(states=states, lens=lens, final=final, final_lens=flen, status=status, fail_step=fail, path_len=d_len)


def _raise_on_failure(out):
    st = out.status.cpu().numpy()
    if st.any():
        p = int(np.flatnonzero(st)[0])
        raise AssertionError("path %d, step %d: the move leaves an invalid presentation (row status %d; ACMove raises here)"
                             % (p, int(out.fail_step[p]), int(st[p])))


def replay_final_state(action_list, initial_state, device="cuda:0"):
    """agents/training.py:74-91"""
    out = replay_paths(np.asarray(initial_state)[None, :], [list(action_list)], True, device, keep_states=False)
    _raise_on_failure(out)
    return out.final[0].cpu().numpy()


def actions_to_path(action_list, initial_state, device="cuda:0"):
    """agents/training.py:94-135: ([[action, total_length], ...], [{"action", "r0", "r1", "lengths"}, ...])"""
    action_list = [int(a) for a in action_list]
    out = replay_paths(np.asarray(initial_state)[None, :], [action_list], True, device)
    _raise_on_failure(out)
    n = len(action_list)
    states = out.states[0, :n].cpu().numpy()
    lens = out.lens[0, :n].cpu().numpy()
    m = states.shape[1] // 2 if n else 0
    path, detailed = [], []
    for t, a in enumerate(action_list):
        path.append([a, int(lens[t, 0]) + int(lens[t, 1])])
        detailed.append({"action": a, "r0": [int(x) for x in states[t, :m] if x != 0], "r1": [int(x) for x in states[t, m:] if x != 0],
                         "lengths": [int(lens[t, 0]), int(lens[t, 1])]})
    return path, detailed
