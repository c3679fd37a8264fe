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


def generate_miller_schupp_presentations(n, max_w_len, _backend=None):
    """dict {len(w): [presentation, ...]}: every presentation is the list relator1 + relator2, zero padded to
    max_relator_length = 2 * max(2n + 3, max_w_len + 1) + 2 each (miller_schupp.py:43-46, 79-80)."""
    assert n >= 1 and max_w_len >= 1, f"expect n >= 1 and max_w_len >=1 ; got n = {n}, max_w_len = {max_w_len}"
    if max_w_len > 15:
        raise ValueError("max_w_len <= 15 (x^-1 w is held in 32 bits)")
    be = _backend if _backend is not None else _CudaMs()
    max_relator_length = 2 * max(2 * n + 3, max_w_len + 1) + 2
    relator1 = [-1] + [2] * n + [1] + [-2] * (n + 1) + [0] * (max_relator_length - 2 * n - 3)
    out = {}
    for packed in _unique_relators(max_w_len, be):
        length = int(packed >> 32)
        letters = [int(_LETTER[(int(packed) >> (2 * i)) & 3]) for i in range(length)]
        out.setdefault(length - 1, []).append(relator1 + letters + [0] * (max_relator_length - length))
    return out


def write_list_to_text_file(list, filepath):
    if not filepath.endswith(".txt"):
        filepath = filepath + ".txt"
    with open(filepath, "w") as f:
        for element in list:
            f.write(f"{element}\n")


def trivialize_miller_schupp_through_search(min_n, max_n, min_w_len, max_w_len, max_nodes_to_explore, search_fn,
                                            write_output_to_file=False):
    """Apply greedy_search or bfs to every MS(n, w) with n in [min_n, max_n], len(w) in [min_w_len, max_w_len]
    (miller_schupp.py:95-172).  Returns (solved_rels, unsolved_rels, solved_paths) in the reference's order."""
    assert search_fn.__name__ in ["greedy_search", "bfs"], f"expect search_fn to be greedy or bfs; got {search_fn.__name__}"
    from acx import search as _search

    greedy = search_fn.__name__ == "greedy_search"
    dev = str(_device.device())
    solved_rels, unsolved_rels, solved_paths = [], [], []
    for n in range(min_n, max_n + 1):
        rels = generate_miller_schupp_presentations(n, max_w_len)
        for lenw in range(min_w_len, max_w_len + 1):
            print(f"Applying {search_fn.__name__} to presentations of n = {n}, lenw = {lenw}")
            batch = rels[lenw]  # KeyError where the reference raises it (no word of that length)
            if not batch:
                continue
            res, paths = _search.search_batch(np.array(batch, dtype=np.int8), max_nodes_to_explore, False, greedy, dev)
            for pres, r, path in zip(batch, res, paths):
                if r[_search.RES_ERROR]:
                    raise AssertionError("a relator became empty during the search: not a valid presentation")
                if r[_search.RES_SOLVED]:
                    solved_rels.append(pres)
                    solved_paths.append(path)
                else:
                    unsolved_rels.append(pres)
    if write_output_to_file:
        dirname = os.path.join(os.path.dirname(os.path.realpath(__file__)), "data")
        os.makedirs(dirname, exist_ok=True)
        filename_base = f"n-{min_n}-to-{max_n}_lenw-{min_w_len}-to-{max_w_len}-max-nodes-{max_nodes_to_explore}-{search_fn.__name__}"
        filepath_base = os.path.join(dirname, filename_base)
        write_list_to_text_file(list=solved_rels, filepath=filepath_base + "_solved")
        write_list_to_text_file(list=unsolved_rels, filepath=filepath_base + "_unsolved")
        write_list_to_text_file(list=solved_paths, filepath=filepath_base + "_paths")
        print(f"""saved output in {dirname} with filenames: 
                {filename_base + "_solved"}
                {filename_base + "_unsolved"}
                {filename_base + "_paths"}""")
    return solved_rels, unsolved_rels, solved_paths


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--min-n", type=int, default=1, help="Minimum value of label n of Miller-Schupp series")
    parser.add_argument("--max-n", type=int, default=7, help="Maximum value of label n of Miller-Schupp series")
    parser.add_argument("--min-w-len", type=int, default=1, help="Minimum word-length of w of Miller-Schupp series")
    parser.add_argument("--max-w-len", type=int, default=7, help="Maximum word-length of w of Miller-Schupp series")
    parser.add_argument("--max-nodes-to-explore", type=int, default=int(1e6), help="Maximum number of nodes to explore")
    parser.add_argument("--search-fn", type=str, default="greedy", help="search function: greedy or bfs")
    args = parser.parse_args()
    from ac_solver.search.breadth_first import bfs
    from ac_solver.search.greedy import greedy_search

    assert args.search_fn in ["greedy", "bfs"], f"expect search-algorithm to be greedy or bfs; got {args.search_fn}"
    assert args.min_n <= args.max_n, "min_n cannot be greater than max_n"
    assert args.min_w_len <= args.max_w_len, "min_w_len cannot be greater than max_w_len"
    trivialize_miller_schupp_through_search(args.min_n, args.max_n, args.min_w_len, args.max_w_len, args.max_nodes_to_explore,
                                            greedy_search if args.search_fn == "greedy" else bfs, write_output_to_file=True)
