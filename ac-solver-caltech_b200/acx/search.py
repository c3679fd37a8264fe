"""Batched greedy_search / bfs on the GPU (C ABI: acx_search_batch, one warp per search instance)."""
import numpy as np
import torch

from . import _lib
from .batch import _prep

RES_SOLVED, RES_SEEN, RES_EXPANDED, RES_PATHLEN, RES_ERROR, RES_LASTNODE, RES_LASTACT, RES_LASTLEN = range(8)


def search_batch(roots, budget, cyclical=False, greedy=True, device="cuda:0", path_cap=4096):
    """roots: array-like [S, 2m] (one width).  Returns (result int32 [S,8], paths: list of list[(a, len)] or None)."""
    roots = np.ascontiguousarray(np.asarray(roots), dtype=np.int8)
    if roots.ndim != 2 or roots.shape[1] % 2:
        raise ValueError("roots must be [S, 2*max_relator_length]")
    S, row = roots.shape
    m = row // 2
    dev = torch.device(device)
    stream = _prep(dev)
    L = _lib.lib()
    need = L.acx_search_workspace_bytes(S, m, int(budget))
    if need < 0:
        _lib.check(int(need), "acx_search_workspace_bytes")
    ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
    d_roots = torch.from_numpy(roots).to(dev)
    while True:
        result = torch.zeros((S, 8), dtype=torch.int32, device=dev)
        path = torch.zeros((S, path_cap, 2), dtype=torch.int32, device=dev)
        rc = L.acx_search_batch(d_roots.data_ptr(), S, m, int(budget), int(bool(cyclical)), int(bool(greedy)),
                                ws.data_ptr(), int(need), result.data_ptr(), path.data_ptr(), path_cap, stream)
        _lib.check(rc, "acx_search_batch")
        res = result.cpu().numpy()
        short = res[:, RES_PATHLEN] < 0
        if not short.any():
            break
        path_cap = int(-res[:, RES_PATHLEN].min()) + 8  # deeper than expected: rerun with room
    plen = res[:, RES_PATHLEN]
    maxlen = int(plen.max()) if S else 0
    hp = path[:, :max(maxlen, 1)].cpu().numpy()
    paths = []
    for s in range(S):
        n = int(plen[s])
        paths.append([(int(a), int(b)) for a, b in hp[s, :n]] if n > 0 else None)
    return res, paths


def search_one(presentation, max_nodes_to_explore, cyclical, greedy, device="cuda:0"):
    row = np.asarray(presentation)
    res, paths = search_batch(row[None, :], max_nodes_to_explore, cyclical, greedy, device)
    if res[0, RES_ERROR]:
        raise AssertionError("a relator became empty during the search: not a valid presentation")
    return bool(res[0, RES_SOLVED]), paths[0], int(res[0, RES_SEEN]), int(res[0, RES_EXPANDED])
