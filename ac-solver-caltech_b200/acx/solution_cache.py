"""Solution memoisation of the value-guided searches and the path post-processing next to it
(value_search/value_guided_search.py:100-250, 563-635), with the reference's names and results:

    expand_path_with_cyclic_reductions(presentation, compact_path)     # one path; expand_paths_... for a batch
    backfill_solution_cache(cache, presentation, path, cyclically_reduce=False, expand_rotations=True)
    _check_cache_with_rotations(state_tup, mrl, solution_cache), _expand_cache_with_rotations(...)

The cache is the reference's: a dict {tuple(state): [(action, total_length), ...]}.  Everything that applies AC moves
runs on the GPU (acx_path_expand_cyclic, acx_path_replay); the rotation bookkeeping is dictionary work on the host.
"""
import numpy as np
import torch

from . import _lib
from .batch import _prep, _ptr
from .replay import replay_paths

# conjugation move r_i -> g r_i g^-1 by (relator index, g)  (value_guided_search.py:119-122, 169-172)
CONJ_MOVE = {(0, 1): 7, (0, -1): 11, (0, 2): 9, (0, -2): 5, (1, 1): 8, (1, -1): 4, (1, 2): 10, (1, -2): 6}


class _CudaExpand:
    def __init__(self, device):
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)

    def expand(self, *args):
        _lib.check(_lib.lib().acx_path_expand_cyclic(*args, self.stream), "acx_path_expand_cyclic")


def expand_paths_with_cyclic_reductions(roots, compact_paths, device="cuda:0", _backend=None):
    """Batch form: roots [P, 2m], compact_paths = P lists of (action, total_length).  Returns P expanded lists."""
    roots = np.ascontiguousarray(np.asarray(roots), dtype=np.int8)
    P, rowb = roots.shape
    m = rowb // 2
    be = _backend if _backend is not None else _CudaExpand(device)
    T = max([len(p) for p in compact_paths], default=0)
    acts = np.zeros((P, max(T, 1)), dtype=np.uint8)
    plen = np.zeros((P,), dtype=np.int32)
    for p, path in enumerate(compact_paths):
        a = np.asarray([int(x[0]) for x in path], dtype=np.int64)
        if a.size and (a.min() < 0 or a.max() > 11):
            raise AssertionError("move ids are 0..11")  # ACMove asserts (ac_moves.py:188-190)
        acts[p, : a.size] = a
        plen[p] = a.size
    d_roots, d_acts, d_len = (torch.from_numpy(x).to(be.dev) for x in (roots, acts, plen))
    cap = max(2 * T + 8, 8)
    while True:
        oa = torch.empty((P, cap), dtype=torch.uint8, device=be.dev)
        ot = torch.empty((P, cap), dtype=torch.int16, device=be.dev)
        olen = torch.empty((P,), dtype=torch.int32, device=be.dev)
        fail = torch.empty((P,), dtype=torch.int32, device=be.dev)
        status = torch.empty((P,), dtype=torch.uint8, device=be.dev)
        be.expand(_ptr(d_roots), _ptr(d_acts), _ptr(d_len), P, acts.shape[1] if T else 0, m, _ptr(oa), _ptr(ot), cap, _ptr(olen),
                  _ptr(fail), _ptr(status))
        st = status.cpu().numpy()
        if (st == 3).any():  # every move can be followed by up to m / 2 conjugations
            cap = min(2 * cap, T * (m // 2 + 1) + 8)
            continue
        break
    if st.any():
        p = int(np.flatnonzero(st)[0])
        if st[p] == 1:
            raise AssertionError("path %d, step %d: a relator becomes empty (ACMove raises, utils.py:264-266)" % (p, int(fail[p])))
        raise KeyError("path %d, step %d: not a presentation over x, y (no conjugation move for its first letter)" % (p, int(fail[p])))
    oa, ot, olen = oa.cpu().numpy(), ot.cpu().numpy(), olen.cpu().numpy()
    return [[(int(a), int(t)) for a, t in zip(oa[p, : olen[p]], ot[p, : olen[p]])] for p in range(P)]


def expand_path_with_cyclic_reductions(presentation, compact_path, device="cuda:0", _backend=None):
    """value_guided_search.py:581-635.  An empty path is returned as it is (:600-601)."""
    if not compact_path:
        return compact_path
    return expand_paths_with_cyclic_reductions(np.asarray(presentation, dtype=np.int8)[None, :], [compact_path], device, _backend)[0]


def _rotations(state, mrl):
    """For each relator whose ends do not cancel: (relator index, letters, offset).  Relators of length <= 1 and relators
    whose first and last letters are inverse are skipped (their rotations are not freely reduced, :131-133, :184-186)."""
    for rel in range(2):
        off = rel * mrl
        nz = state[off:off + mrl]
        nz = nz[nz != 0]
        if len(nz) > 1 and int(nz[-1]) + int(nz[0]) != 0:
            yield rel, nz, off


def _check_cache_with_rotations(state_tup, mrl, solution_cache):
    """Path to trivial for this state or for a cyclic rotation of one of its relators (prefixed with the conjugations
    that rotate it), else None (:100-151)."""
    if state_tup in solution_cache:
        return list(solution_cache[state_tup])
    state = np.array(state_tup, dtype=np.int8)
    total = int(np.count_nonzero(state))
    for rel, nz, off in _rotations(state, mrl):
        moves = []
        for k in range(1, len(nz)):
            key = (rel, -int(nz[k - 1]))  # rotating a_k to the back = conjugating by a_k^-1 (:137-139)
            if key not in CONJ_MOVE:
                break
            moves.append((CONJ_MOVE[key], total))
            rotated = state.copy()
            rotated[off:off + len(nz)] = np.concatenate([nz[k:], nz[:k]])
            hit = solution_cache.get(tuple(rotated))
            if hit is not None:
                return list(moves) + list(hit)
    return None


def _expand_cache_with_rotations(cache, state_tup, suffix_path, mrl, cyclically_reduce):
    """Insert every cyclic rotation of both relators with the conjugations that undo it in front of the suffix
    (:154-209).  Existing entries are kept."""
    state = np.array(state_tup, dtype=np.int8)
    total = int(np.count_nonzero(state))
    for rel, nz, off in _rotations(state, mrl):
        undo = []
        for k in range(1, len(nz)):
            key = (rel, int(nz[k - 1]))  # undo: conjugate by a_k, then a_{k-1}, ..., a_1 (:191-199)
            if key not in CONJ_MOVE:
                break
            undo.append((CONJ_MOVE[key], total))
            rotated = state.copy()
            rotated[off:off + len(nz)] = np.concatenate([nz[k:], nz[:k]])
            tup = tuple(rotated)
            if tup not in cache:
                cache[tup] = list(reversed(undo)) + suffix_path


def _replay_states(root, actions, cyclical, device):
    out = replay_paths(root[None, :], [actions], cyclical=cyclical, device=device)
    st = int(out.status[0])
    if st:  # the reference applies every move of the path, the last one included (:246-247)
        raise AssertionError("step %d of the path leaves an invalid presentation (row status %d)" % (int(out.fail_step[0]), st))
    return out.states[0, : len(actions)].cpu().numpy()


def backfill_solution_cache(cache, presentation, path, cyclically_reduce=False, expand_rotations=True, device="cuda:0",
                            _replay=_replay_states):
    """Replay a solved path and cache the remaining suffix at every intermediate state (:212-250).  The replay is one
    acx_path_replay launch; a move that empties a relator raises AssertionError like ACMove."""
    root = np.array(presentation, dtype=np.int8)
    mrl = len(root) // 2
    path = list(path)
    if not path:
        return cache
    states = [root] + list(_replay(root, [int(a) for a, _ in path], cyclically_reduce, device)[: len(path) - 1])
    for i, state in enumerate(states):
        tup = tuple(int(v) for v in state)
        suffix = list(path[i:])
        if tup not in cache:
            cache[tup] = suffix
        if expand_rotations:
            _expand_cache_with_rotations(cache, tup, suffix, mrl, cyclically_reduce)
    return cache
