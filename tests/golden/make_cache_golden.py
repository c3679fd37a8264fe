"""Generates tests/golden/cache_golden.json by RUNNING THE REFERENCE (value_search/value_guided_search.py imported from
/root/reference; build container only): compact paths of the length-priority search with cyclical reduction and their
expansion by expand_path_with_cyclic_reductions, the dictionaries backfill_solution_cache builds (with and without
rotation variants), _check_cache_with_rotations lookups of rotated states, and value_guided_greedy_search runs that end
in a cache hit.

    python tests/golden/make_cache_golden.py
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_loader  # noqa: E402
from helpers import dataset_rows  # noqa: E402


def cache_to_list(cache):
    return sorted([[int(v) for v in k], [[int(a), int(t)] for a, t in p]] for k, p in cache.items())


def main():
    ref_loader._install_gymnasium_stand_in()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    try:
        vgs = importlib.import_module("value_search.value_guided_search")
        rows = dataset_rows(HERE)
        ak2 = np.array([1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0], dtype=np.int8)
        pool = [ak2] + [rows[i] for i in (0, 1, 2, 3, 5, 8, 13, 21, 34, 55, 89, 144, 170, 200, 233, 340, 377, 400, 510, 520)]
        out = {"expand": [], "backfill": [], "check": [], "search": []}
        solved = []
        for pres in pool:
            for cyc in (True, False):
                ok, path, _ = vgs.value_guided_greedy_search(pres, use_length_priority=True, cyclically_reduce_after_moves=cyc,
                                                             max_nodes_to_explore=3000)
                if not ok:
                    continue
                path = [(int(a), int(t)) for a, t in path]
                # a path found WITHOUT cyclic reduction is expanded too: the inserted conjugations change the states the
                # later moves act on, so the result is whatever the reference computes (or the exception it raises)
                try:
                    expanded = [[int(a), int(t)] for a, t in vgs.expand_path_with_cyclic_reductions(pres, path)]
                except (AssertionError, KeyError) as e:
                    expanded = type(e).__name__
                out["expand"].append({"presentation": pres.tolist(), "cyclical": cyc, "compact": path, "expanded": expanded})
                solved.append((pres, cyc, path))
        for pres, cyc, path in solved[:12]:
            for rot in (False, True):
                cache = {}
                vgs.backfill_solution_cache(cache, pres, path, cyclically_reduce=cyc, expand_rotations=rot)
                out["backfill"].append({"presentation": pres.tolist(), "cyclical": cyc, "path": path, "expand_rotations": rot,
                                        "cache": cache_to_list(cache)})
        # lookups: a cache WITHOUT rotation variants, queried with rotated states (test_rotation_cache.py, test 3)
        cache = {}
        for pres, cyc, path in solved:
            if not cyc:
                vgs.backfill_solution_cache(cache, pres, path, cyclically_reduce=False, expand_rotations=False)
        rng = np.random.default_rng(11)
        keys = sorted(cache.keys())
        for idx in rng.choice(len(keys), size=min(60, len(keys)), replace=False):
            state = np.array(keys[idx], dtype=np.int8)
            mrl = len(state) // 2
            rel = int(rng.integers(0, 2))
            nz = state[rel * mrl:(rel + 1) * mrl]
            nz = nz[nz != 0]
            if len(nz) < 2:
                continue
            k = int(rng.integers(1, len(nz)))
            q = state.copy()
            q[rel * mrl: rel * mrl + len(nz)] = np.concatenate([nz[-k:], nz[:-k]])
            small = {kk: vv for kk, vv in cache.items() if len(kk) == len(q)}
            res = vgs._check_cache_with_rotations(tuple(q), mrl, small)
            out["check"].append({"query": q.tolist(), "result": None if res is None else [[int(a), int(t)] for a, t in res]})
        out["check_cache"] = cache_to_list(cache)
        # searches that stop on a cache hit: cache filled from the first half of the pool, searched on all of it
        cache = {}
        for pres, cyc, path in solved[: len(solved) // 2]:
            if not cyc:
                vgs.backfill_solution_cache(cache, pres, path, cyclically_reduce=False, expand_rotations=True)
        out["search_cache"] = cache_to_list(cache)
        for pres in pool:
            sub = {k: v for k, v in cache.items() if len(k) == len(pres)}
            ok, path, stats = vgs.value_guided_greedy_search(pres, use_length_priority=True, max_nodes_to_explore=3000,
                                                             solution_cache=sub)
            out["search"].append({"presentation": pres.tolist(), "solved": bool(ok), "path": [[int(a), int(t)] for a, t in path],
                                  "stats": {k: (bool(v) if isinstance(v, bool) else int(v)) for k, v in stats.items()}})
    finally:
        sys.path.remove(ref_loader.REFERENCE_ROOT)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.") or k.startswith("value_search")]:
            sys.modules.pop(k)
        sys.modules.update(saved)
    with open(os.path.join(HERE, "cache_golden.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print({k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
