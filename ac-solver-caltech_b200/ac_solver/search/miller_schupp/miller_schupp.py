"""Miller-Schupp presentations MS(n, w) = <x, y | x^-1 y^n x = y^(n+1), x = w> and the search sweep over them,
with the reference's names, arguments and results (ac_solver/search/miller_schupp/miller_schupp.py).

    python -u -m ac_solver.search.miller_schupp.miller_schupp --min-n=3 --max-n=5 --min-w-len=2 --max-w-len=3 --search-fn=greedy

The 4^k words of every length are enumerated, filtered and reduced by one CUDA launch per length
(acx_ms_enumerate); the `seen` set of the reference becomes a stable sort on the class key.  The sweep hands all
presentations of one width to the batched search engine (one warp per instance) instead of one Python search
at a time.  There is no CPU path.
"""
import argparse
import os

import numpy as np
import torch

from ac_solver import _device
from acx import _lib
from acx.batch import _prep

_LETTER = np.array([1, 2, -1, -2], dtype=np.int64)
_CHUNK = 1 << 24


class _CudaMs:
    def __init__(self):
        self.dev = _device.device()
        self.stream = _prep(self.dev)

    def enumerate(self, w_len, first, count, key, rel):
        _lib.check(_lib.lib().acx_ms_enumerate(w_len, first, count, key.data_ptr(), rel.data_ptr(), self.stream),
                   "acx_ms_enumerate")


def _unique_relators(max_w_len, be):
    """Reduced x^-1 w of the first word of every cyclic-permutation class, in the reference's enumeration order
    (search_len ascending, itertools.product order inside, miller_schupp.py:50-75).  Returns int64 packed words."""
    keys, rels = [], []
    for k in range(1, max_w_len + 1):
        total = 4 ** k
        for first in range(0, total, _CHUNK):
            cnt = min(_CHUNK, total - first)
            key = torch.empty((cnt,), dtype=torch.int64, device=be.dev)
            rel = torch.empty((cnt,), dtype=torch.int64, device=be.dev)
            be.enumerate(k, first, cnt, key, rel)
            ok = key != -1  # ~0: filtered out
            keys.append(key[ok])
            rels.append(rel[ok])
    key, rel = torch.cat(keys), torch.cat(rels)
    if key.numel() == 0:
        return np.zeros((0,), dtype=np.int64)
    skey, perm = torch.sort(key, stable=True)  # equal keys keep enumeration order: the first one is the word kept
    first_of_run = torch.ones_like(skey, dtype=torch.bool)
    first_of_run[1:] = skey[1:] != skey[:-1]
    winners = torch.sort(perm[first_of_run]).values
    return rel[winners].cpu().numpy()


def _presentations_from_relators(n, max_w_len, packed_relators):
    """{len(w): [relator1 + relator2, ...]} for one n from the packed second relators (which do not depend on n)"""
    max_relator_length = 2 * max(2 * n + 3, max_w_len + 1) + 2
    relator1 = [-1] + [2] * n + [1] + [-2] * (n + 1) + [0] * (max_relator_length - 2 * n - 3)
    out = {}
    for packed in packed_relators:
        length = int(packed >> 32)
        letters = [int(_LETTER[(int(packed) >> (2 * i)) & 3]) for i in range(length)]
        out.setdefault(length - 1, []).append(relator1 + letters + [0] * (max_relator_length - length))
    return out


def generate_miller_schupp_presentations(n, max_w_len, _backend=None):
    """dict {len(w): [presentation, ...]}: every presentation is the list relator1 + relator2, zero padded to
    max_relator_length = 2 * max(2n + 3, max_w_len + 1) + 2 each (miller_schupp.py:43-46, 79-80)."""
    assert n >= 1 and max_w_len >= 1, f"expect n >= 1 and max_w_len >=1 ; got n = {n}, max_w_len = {max_w_len}"
    if max_w_len > 15:
        raise ValueError("max_w_len <= 15 (x^-1 w is held in 32 bits)")
    be = _backend if _backend is not None else _CudaMs()
    return _presentations_from_relators(n, max_w_len, _unique_relators(max_w_len, be))


def write_list_to_text_file(list, filepath):
    """one element per line; `.txt` is appended to the path when missing (same helper name and arguments as the reference)"""
    path = filepath if filepath.endswith(".txt") else filepath + ".txt"
    with open(path, "w") as out:
        out.writelines(str(item) + "\n" for item in list)


def _search_in_chunks(batch, budget, greedy, dev):
    """search_batch over `batch`, split so that the per-instance workspaces (about 112 MB each at a 1e6-node budget)
    stay inside a fraction of the free device memory"""
    from acx import search as _search

    rows = np.array(batch, dtype=np.int8)
    per = int(_lib.lib().acx_search_workspace_bytes(1, rows.shape[1] // 2, int(budget)))
    free = torch.cuda.mem_get_info(torch.device(dev))[0] if torch.cuda.is_available() else 1 << 34
    step = max(1, int(0.6 * free // max(per, 1)))
    results, paths = [], []
    for a in range(0, len(rows), step):
        res, pth = _search.search_batch(rows[a:a + step], budget, False, greedy, dev)
        results.extend(res)
        paths.extend(pth)
    return results, paths


def trivialize_miller_schupp_through_search(min_n, max_n, min_w_len, max_w_len, max_nodes_to_explore, search_fn,
                                            write_output_to_file=False):
    """Apply greedy_search or bfs to every MS(n, w) with n in [min_n, max_n], len(w) in [min_w_len, max_w_len]
    (miller_schupp.py:95-172).  Returns (solved_rels, unsolved_rels, solved_paths) in the reference's order."""
    assert search_fn.__name__ in ["greedy_search", "bfs"], f"expect search_fn to be greedy or bfs; got {search_fn.__name__}"
    from acx import search as _search

    greedy = search_fn.__name__ == "greedy_search"
    dev = str(_device.device())
    solved_rels, unsolved_rels, solved_paths = [], [], []
    assert max_w_len >= 1, f"expect max_w_len >= 1; got {max_w_len}"
    if max_w_len > 15:
        raise ValueError("max_w_len <= 15 (x^-1 w is held in 32 bits)")
    packed = _unique_relators(max_w_len, _CudaMs())  # the 4^k enumeration does not depend on n: once per sweep
    for n in range(min_n, max_n + 1):
        assert n >= 1, f"expect n >= 1; got n = {n}"
        rels = _presentations_from_relators(n, max_w_len, packed)
        for lenw in range(min_w_len, max_w_len + 1):
            print(f"Applying {search_fn.__name__} to presentations of n = {n}, lenw = {lenw}")
            batch = rels[lenw]  # KeyError where the reference raises it (no word of that length)
            if not batch:
                continue
            res, paths = _search_in_chunks(batch, max_nodes_to_explore, greedy, dev)
            for pres, r, path in zip(batch, res, paths):
                if r[_search.RES_ERROR]:
                    raise AssertionError("a relator became empty during the search: not a valid presentation")
                if r[_search.RES_SOLVED]:
                    solved_rels.append(pres)
                    solved_paths.append(path)
                else:
                    unsolved_rels.append(pres)
    if write_output_to_file:
        # same directory and file names as the reference writes (data/ next to this module, three lists per sweep)
        data_dir = os.path.join(os.path.dirname(os.path.realpath(__file__)), "data")
        os.makedirs(data_dir, exist_ok=True)
        stem = "n-%d-to-%d_lenw-%d-to-%d-max-nodes-%d-%s" % (min_n, max_n, min_w_len, max_w_len, max_nodes_to_explore, search_fn.__name__)
        written = []
        for suffix, items in (("_solved", solved_rels), ("_unsolved", unsolved_rels), ("_paths", solved_paths)):
            write_list_to_text_file(items, os.path.join(data_dir, stem + suffix))
            written.append(stem + suffix)
        print("saved output in %s with filenames: %s" % (data_dir, ", ".join(written)))
    return solved_rels, unsolved_rels, solved_paths


def _cli():
    """same flags and defaults as the reference's command line"""
    ap = argparse.ArgumentParser(description="Trivialise Miller-Schupp presentations MS(n, w) by greedy search or bfs on the GPU")
    for flag, default, what in (("--min-n", 1, "smallest n"), ("--max-n", 7, "largest n"), ("--min-w-len", 1, "shortest w"),
                                ("--max-w-len", 7, "longest w"), ("--max-nodes-to-explore", int(1e6), "node budget per search")):
        ap.add_argument(flag, type=int, default=default, help=what)
    ap.add_argument("--search-fn", type=str, default="greedy", help="greedy or bfs")
    a = ap.parse_args()
    from ac_solver.search.breadth_first import bfs
    from ac_solver.search.greedy import greedy_search

    assert a.search_fn in ["greedy", "bfs"], f"expect search-algorithm to be greedy or bfs; got {a.search_fn}"
    assert a.min_n <= a.max_n, "min_n cannot be greater than max_n"
    assert a.min_w_len <= a.max_w_len, "min_w_len cannot be greater than max_w_len"
    trivialize_miller_schupp_through_search(a.min_n, a.max_n, a.min_w_len, a.max_w_len, a.max_nodes_to_explore,
                                            greedy_search if a.search_fn == "greedy" else bfs, write_output_to_file=True)


if __name__ == "__main__":
    _cli()
