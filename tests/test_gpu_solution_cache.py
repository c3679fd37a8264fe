"""GPU parity of acx.solution_cache and of the cache-aware searches with vectors produced by the reference
(tests/golden/cache_golden.json): expand_path_with_cyclic_reductions, backfill_solution_cache (GPU replay), and
value_guided_greedy_search / beam_search runs that end in a cache hit."""
import json
import os

import numpy as np
import pytest

from test_solution_cache import cache_to_list, to_cache

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "cache_golden.json")) as f:
        return json.load(f)


def test_expand_path_with_cyclic_reductions(cuda, golden, oracle):
    from acx.solution_cache import expand_path_with_cyclic_reductions

    assert expand_path_with_cyclic_reductions(golden["expand"][0]["presentation"], []) == []
    for case in golden["expand"]:
        pres = np.array(case["presentation"], dtype=np.int8)
        compact = [tuple(x) for x in case["compact"]]
        if isinstance(case["expanded"], str):  # the reference raised
            with pytest.raises((AssertionError, KeyError)):
                expand_path_with_cyclic_reductions(pres, compact)
            continue
        got = expand_path_with_cyclic_reductions(pres, compact)
        assert [list(x) for x in got] == case["expanded"]
        if not case["cyclical"]:
            continue
        # a path found with cyclic reduction, once expanded, replays to the trivial presentation under free reduction only
        state, m = pres.copy(), len(pres) // 2
        lens = [int(np.count_nonzero(state[:m])), int(np.count_nonzero(state[m:]))]
        for a, t in got:
            state, lens = oracle.ACMove(a, state, m, lens, False)
            assert sum(lens) == t
        assert sum(lens) == 2


def test_backfill_solution_cache(cuda, golden):
    from acx.solution_cache import backfill_solution_cache

    for case in golden["backfill"]:
        cache = {}
        backfill_solution_cache(cache, case["presentation"], [tuple(x) for x in case["path"]], cyclically_reduce=case["cyclical"],
                                expand_rotations=case["expand_rotations"])
        assert cache_to_list(cache) == case["cache"]


def test_value_guided_greedy_search_with_cache(cuda, golden):
    from acx.beam import value_guided_greedy_search

    cache = to_cache(golden["search_cache"])
    hits = 0
    for case in golden["search"]:
        pres = np.array(case["presentation"], dtype=np.int8)
        sub = {k: v for k, v in cache.items() if len(k) == len(pres)}
        ok, path, stats = value_guided_greedy_search(pres, use_length_priority=True, max_nodes_to_explore=3000, solution_cache=sub,
                                                     device=str(cuda))
        assert ok == case["solved"]
        assert [list(x) for x in path] == case["path"]
        assert {k: int(v) for k, v in stats.items()} == {k: int(v) for k, v in case["stats"].items()}
        hits += int(bool(stats.get("cache_hit")))
    assert hits >= 10


def test_beam_search_with_cache(cuda, golden, oracle):
    import torch

    from acx.beam import beam_search
    from test_gpu_beam import _beam_oracle

    model = torch.nn.Linear(14, 1).to(cuda).eval()  # score = total length
    with torch.no_grad():
        model.weight.zero_()
        model.bias.zero_()
        model.weight[0, 0] = 1.0
    mean, std = np.zeros(14, np.float32), np.ones(14, np.float32)
    cache = to_cache(golden["search_cache"])
    hits = 0
    for case in golden["search"]:
        pres = np.array(case["presentation"], dtype=np.int8)
        sub = {k: v for k, v in cache.items() if len(k) == len(pres)}
        for width, budget in ((8, 1500), (1, 60)):
            want = _beam_oracle(oracle, pres, model, "mlp", mean, std, width, budget, cuda, False, cache=sub)
            got = beam_search(pres, model, architecture="mlp", feat_mean=mean, feat_std=std, beam_width=width,
                              max_nodes_to_explore=budget, device=str(cuda), solution_cache=sub)
            assert got == want, (len(pres), width, budget)
            hits += int(bool(want[2].get("cache_hit")))
    assert hits >= 5
