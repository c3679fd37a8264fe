"""GPU parity of the Miller-Schupp drop-in (acx_ms_enumerate + the batched search sweep) with the
reference-generated fixtures."""
import json
import os

import numpy as np
import pytest

from test_host_ms import check_against_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ms(cuda):
    from ac_solver.search.miller_schupp import miller_schupp

    return miller_schupp


def test_generator_matches_the_reference(ms, golden_dir):
    with open(os.path.join(golden_dir, "ms_golden.json")) as f:
        golden = json.load(f)
    check_against_golden(ms, golden, None)


def test_long_words_keep_one_word_per_rotation_class(ms):
    d = ms.generate_miller_schupp_presentations(1, 11)
    seen = set()
    for lenw, ps in d.items():
        for p in ps:
            h = len(p) // 2
            r = tuple(v for v in p[h:] if v != 0)
            assert len(r) == lenw + 1 and sum(v for v in r if abs(v) == 1) == -1
            assert not any(r[i] == -r[i + 1] for i in range(len(r) - 1)) and r[0] != -r[-1]
            rots = {r[i:] + r[:i] for i in range(len(r))}
            assert not (rots & seen)
            seen |= rots


def test_sweep_equals_single_searches(ms, golden_dir):
    """trivialize_miller_schupp_through_search (miller_schupp.py:95-172) = greedy_search per presentation, in order;
    the solved sets at max_nodes_to_explore=10000 are pinned by the reference run in greedy10k.json."""
    from ac_solver.search.greedy import greedy_search

    solved, unsolved, paths = ms.trivialize_miller_schupp_through_search(1, 2, 1, 5, 2000, greedy_search)
    want_s, want_u, want_p = [], [], []
    for n in (1, 2):
        d = ms.generate_miller_schupp_presentations(n, 5)
        for lenw in range(1, 6):
            for p in d[lenw]:
                ok, path = greedy_search(p, max_nodes_to_explore=2000)
                (want_s if ok else want_u).append(p)
                if ok:
                    want_p.append(path)
    assert (solved, unsolved, paths) == (want_s, want_u, want_p)
    assert len(solved) + len(unsolved) == 60
