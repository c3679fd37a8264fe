"""oracle/barcode_oracle.c against vectors produced by the reference's own C++ programs
(tests/golden/barcode_golden.json, made by tests/golden/make_barcode_golden.py) and -- where the build
container's oracle/_ref binaries exist -- against those programs run live."""
import hashlib
import json
import os

import numpy as np
import pytest

import barcode_oracle as bo


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "barcode_golden.json")) as f:
        return json.load(f)


def oracle_files(g):
    """The four ac_bfs.cpp files from oracle arrays (independent of the product's writer)."""
    n = len(g["sizes"])
    return {
        "zero_simplices": ('{"0-simplices":[' + "".join("[%d]," % i for i in range(n)) + "[]]}").encode(),
        "zero_filtrations": ('{"0-filt":[' + "".join("%d," % s for s in g["sizes"]) + "-5]}").encode(),
        "one_simplices": ('{"1-simplices":[' + "".join("[%d,%d]," % (a, b) for a, b in g["edges"]) + "[]]}").encode(),
        "one_filtrations": ('{"1-filt":[' + "".join("%d," % s for s in g["efilt"]) + "-5]}").encode(),
    }


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_graph_bfs_matches_reference_files(golden, moves):
    for n in range(2, 11):
        ref = golden["ac_bfs"]["%s_%d" % (moves, n)]
        g = bo.graph_bfs(n, moves == "classic")
        assert (len(g["sizes"]), len(g["edges"])) == (ref["n_nodes"], ref["n_edges"])
        for k, v in oracle_files(g).items():
            assert hashlib.sha256(v).hexdigest() == ref[k], (moves, n, k)
        if "sizes" in ref:
            assert g["sizes"].tolist() == ref["sizes"]
            assert g["edges"].tolist() == ref["edges"]
            assert g["efilt"].tolist() == ref["efilt"]


def test_readme_known_answers(golden):
    # barcode_analysis/5_steps_neibourhoods/README.txt:33-37 (prime moves, radius 5)
    rows = golden["neighbourhoods"]["rows"][:5]
    assert [bo.neighbourhood(r, 5, False) for r in rows] == [28631, 49668, 72392, 28631, 28631]


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_neighbourhood_sizes_match_reference(golden, moves):
    rows, want = golden["neighbourhoods"]["rows"], golden["neighbourhoods"][moves]
    for r, w in list(zip(rows, want))[::3]:
        assert bo.neighbourhood(r, 5, moves == "classic") == w


def test_move_is_sorted_and_reduced():
    rng = np.random.default_rng(5)
    row = np.zeros(16, np.int8)
    row[:3] = [1, 2, 1]
    row[8:9] = [2]
    for _ in range(200):
        t = int(rng.integers(0, 14))
        r = bo.move(row, t, True)
        if r is None:
            continue
        row, size = r
        a, b = row[:8][row[:8] != 0], row[8:][row[8:] != 0]
        assert size == len(a) + len(b)
        assert (len(a), a.tolist()) <= (len(b), b.tolist())
        for w in (a, b):
            assert not np.any(w[:-1] == -w[1:])


@pytest.mark.skipif(not bo.have_ref(), reason="oracle/_ref is only built where /root/reference exists")
def test_against_the_compiled_reference_live():
    for classic in (True, False):
        ref = bo.ref_ac_bfs_files(7, classic)
        assert oracle_files(bo.graph_bfs(7, classic)) == ref
    rows = [[1, 2, 1, -2, 0, 0, 2, 1, -2, 0, 0, 0], [-1, 2, 2, 0, 1, 2, 0, 0]]
    assert bo.ref_neighbourhoods(rows, 4, True) == [bo.neighbourhood(r, 4, True) for r in rows]
