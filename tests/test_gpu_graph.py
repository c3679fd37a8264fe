"""GPU parity of the graph engine (acx.graph over libacx.so's acx_graph_* entry points) with the oracle
(oracle/barcode_oracle.c, pinned to the reference's C++ programs) and with the reference-generated goldens:
node names, sizes, edges and filtrations of ac_bfs.cpp bit for bit, the four output files byte for byte,
ball sizes of neibourhoods.cpp."""
import hashlib
import json
import os

import numpy as np
import pytest

import barcode_oracle as bo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "barcode_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def graph(cuda):
    from acx import graph as g

    return g


def same_graph(r, g):
    return (np.array_equal(r["sizes"], g["sizes"].astype(np.int16)) and np.array_equal(r["edges"], g["edges"])
            and np.array_equal(r["efilt"], g["efilt"].astype(np.int16)))


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_graph_bfs_equals_oracle(graph, moves):
    for n, chunk in ((1, 8), (2, 8), (3, 5), (6, 64), (8, 1000), (10, 1 << 18), (11, 3000)):
        r = graph.ac_graph_bfs(n, moves, chunk=chunk, with_rows=True)
        g = bo.graph_bfs(n, moves == "classic")
        assert same_graph(r, g), (moves, n)
        assert np.array_equal(r["rows"], g["rows"])


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_files_are_byte_identical_to_the_reference(graph, golden, moves, tmp_path):
    for n in (2, 5, 9, 11):
        r = graph.ac_graph_bfs(n, moves)
        ref = golden["ac_bfs"]["%s_%d" % (moves, n)]
        assert (len(r["sizes"]), len(r["edges"])) == (ref["n_nodes"], ref["n_edges"])
        graph.write_simplex_files(r, n, str(tmp_path))
        for k in ("zero_simplices", "zero_filtrations", "one_simplices", "one_filtrations"):
            data = (tmp_path / ("%s_%d" % (k, n))).read_bytes()
            assert hashlib.sha256(data).hexdigest() == ref[k], (moves, n, k)


def test_n12_classic_matches_the_reference_run(graph):
    # the compiled reference at n=12: 791140 nodes, 2293951 edges (31.7 s); the oracle reproduces it in seconds
    r = graph.ac_graph_bfs(12, "classic")
    assert (len(r["sizes"]), len(r["edges"])) == (791140, 2293951)
    assert same_graph(r, bo.graph_bfs(12, True))


@pytest.mark.parametrize("width", [2, 4, 8])
def test_plane_widths_agree(graph, golden, width):
    r = graph.ac_graph_bfs(8, "prime", chunk=700, width_words=width)
    for k, v in graph.simplex_file_contents(r).items():
        assert hashlib.sha256(v).hexdigest() == golden["ac_bfs"]["prime_8"][k]


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_neighbourhood_sizes_match_the_reference(graph, golden, moves):
    rows, want = golden["neighbourhoods"]["rows"], golden["neighbourhoods"][moves]
    got = graph.neighbourhood_sizes(rows, 5, moves == "classic")
    assert got.tolist() == want
    # small batches and chunks: same answer (instance ids restart per batch, steps split the frontier)
    got = graph.neighbourhood_sizes(rows[:9], 5, moves == "classic", instances_per_batch=4, chunk=5000)
    assert got.tolist() == want[:9]


def test_neighbourhood_radii_equal_oracle(graph, golden):
    rows = golden["neighbourhoods"]["rows"][5:9]
    for radius in (0, 1, 2, 4):
        got = graph.neighbourhood_sizes(rows, radius, True)
        assert got.tolist() == [bo.neighbourhood(r, radius, True) for r in rows]


def test_read_do_and_write(graph, tmp_path, capsys):
    src, out = tmp_path / "in.txt", tmp_path / "out.txt"
    src.write_text("[1, 0, 0, 0, 2, 0, 0, 0]\n[1, 2, 1, 0, 2, 0, 0, 0]\n[1, 2, 1,-2, 0, 0, 2, 1, -2, 0, 0, 0]\n[1, 2]\n[2, 1]\n")
    graph.read_do_and_write(str(src), str(out), 5, False)
    assert [int(x) for x in out.read_text().split()] == [28631, 49668, 72392, 28631, 28631]  # README.txt:33-37
    assert capsys.readouterr().out.split("\n")[:5] == ["(a, b)", "(b, aba)", "(baB, abaB)", "(a, b)", "(a, b)"]


def test_domain_errors(graph):
    with pytest.raises(ValueError):
        graph.neighbourhood_sizes([[1, -1, 2, 0]], 2, True)
    long = [1, 2] * 12 + [2, 1] * 12
    with pytest.raises(OverflowError):
        graph.neighbourhood_sizes([long], 3, True, width_words=2)
    assert graph.neighbourhood_sizes([long], 2, True).tolist() == [bo.neighbourhood(long, 2, True)]
