"""acx.solution_cache against vectors produced by the reference (tests/golden/cache_golden.json, made by
tests/golden/make_cache_golden.py): the rotation-aware cache lookups and insertions (host-side dictionary logic), and --
with the product's path code compiled for the CPU (tests/host_harness) -- expand_path_with_cyclic_reductions.
backfill_solution_cache is driven here by the oracle's ACMove for the replay (the GPU replay is covered by the gpu suite)."""
import json
import os

import numpy as np
import pytest

from host_harness import harness


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "cache_golden.json")) as f:
        return json.load(f)


def to_cache(lst):
    return {tuple(k): [tuple(x) for x in p] for k, p in lst}


def cache_to_list(cache):
    return sorted([[int(v) for v in k], [[int(a), int(t)] for a, t in p]] for k, p in cache.items())


def test_check_cache_with_rotations(golden):
    from acx.solution_cache import _check_cache_with_rotations

    cache = to_cache(golden["check_cache"])
    for case in golden["check"]:
        q = case["query"]
        sub = {k: v for k, v in cache.items() if len(k) == len(q)}
        got = _check_cache_with_rotations(tuple(q), len(q) // 2, sub)
        want = case["result"]
        assert (None if got is None else [list(x) for x in got]) == want
    assert _check_cache_with_rotations(tuple(golden["check"][0]["query"]), len(golden["check"][0]["query"]) // 2, {}) is None


def test_expand_path_with_cyclic_reductions_host(golden):
    from acx.solution_cache import expand_path_with_cyclic_reductions, expand_paths_with_cyclic_reductions

    be = harness.HostExpandBackend()
    assert expand_path_with_cyclic_reductions(golden["expand"][0]["presentation"], [], _backend=be) == []
    longer = 0
    for case in golden["expand"]:
        compact = [tuple(x) for x in case["compact"]]
        if isinstance(case["expanded"], str):  # the reference raised
            with pytest.raises((AssertionError, KeyError)):
                expand_path_with_cyclic_reductions(case["presentation"], compact, _backend=be)
            continue
        got = expand_path_with_cyclic_reductions(case["presentation"], compact, _backend=be)
        assert [list(x) for x in got] == case["expanded"]
        longer += len(got) > len(case["compact"])
    assert longer >= 10
    # batch form, ragged lengths, one width
    same = [c for c in golden["expand"] if not isinstance(c["expanded"], str) and len(c["presentation"]) == 36]
    got = expand_paths_with_cyclic_reductions(np.array([c["presentation"] for c in same], dtype=np.int8),
                                              [[tuple(x) for x in c["compact"]] for c in same], _backend=be)
    assert [[list(x) for x in g] for g in got] == [c["expanded"] for c in same]


def test_backfill_solution_cache_with_oracle_replay(golden, oracle):
    from acx.solution_cache import backfill_solution_cache

    def replay(root, actions, cyclical, device):
        state, m = root.copy(), len(root) // 2
        lens = [int(np.count_nonzero(state[:m])), int(np.count_nonzero(state[m:]))]
        out = []
        for a in actions:
            state, lens = oracle.ACMove(a, state, m, lens, cyclical)
            out.append(state.copy())
        return np.stack(out)

    for case in golden["backfill"]:
        cache = {}
        backfill_solution_cache(cache, case["presentation"], [tuple(x) for x in case["path"]], cyclically_reduce=case["cyclical"],
                                expand_rotations=case["expand_rotations"], _replay=replay)
        assert cache_to_list(cache) == case["cache"]
