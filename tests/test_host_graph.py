"""The product's graph engine (csrc/acx_graph.cuh: functors + graph_step<>) compiled for the CPU
(tests/host_harness) and driven by the product's own acx.graph host code, against the oracle and the
reference-generated goldens.  Candidate order is permuted (order 1, 2) to exercise the claim-replacement
path of the visited-set probe that concurrent GPU threads take."""
import hashlib
import json
import os

import numpy as np
import pytest

import barcode_oracle as bo
from acx import graph
from host_harness import harness


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "barcode_golden.json")) as f:
        return json.load(f)


def same_graph(r, g):
    return (np.array_equal(r["sizes"], g["sizes"].astype(np.int16)) and np.array_equal(r["edges"], g["edges"])
            and np.array_equal(r["efilt"], g["efilt"].astype(np.int16)))


@pytest.mark.parametrize("order", [0, 1, 2])
@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_graph_bfs_equals_oracle(order, moves):
    for n, chunk in ((1, 8), (2, 8), (3, 5), (6, 64), (8, 1000)):
        r = graph.ac_graph_bfs(n, moves, chunk=chunk, with_rows=True, _backend=harness.HostGraphBackend(order))
        g = bo.graph_bfs(n, moves == "classic")
        assert same_graph(r, g), (moves, n)
        assert np.array_equal(r["rows"], g["rows"])


@pytest.mark.parametrize("width", [2, 4, 8])
def test_plane_widths_agree(width, golden):
    r = graph.ac_graph_bfs(7, "classic", chunk=500, width_words=width, _backend=harness.HostGraphBackend(2))
    files = graph.simplex_file_contents(r)
    for k, v in files.items():
        assert hashlib.sha256(v).hexdigest() == golden["ac_bfs"]["classic_7"][k]


def test_files_are_byte_identical_to_the_reference(golden):
    for moves, n in (("classic", 9), ("prime", 9), ("prime", 2), ("classic", 2)):
        r = graph.ac_graph_bfs(n, moves, chunk=4096, _backend=harness.HostGraphBackend(1))
        for k, v in graph.simplex_file_contents(r).items():
            assert hashlib.sha256(v).hexdigest() == golden["ac_bfs"]["%s_%d" % (moves, n)][k], (moves, n, k)


@pytest.mark.parametrize("moves", ["classic", "prime"])
def test_neighbourhood_sizes(golden, moves):
    rows, want = golden["neighbourhoods"]["rows"], golden["neighbourhoods"][moves]
    pick = list(range(0, 8)) + [20, 35]
    got = graph.neighbourhood_sizes([rows[i] for i in pick], 5, moves == "classic", instances_per_batch=4, chunk=20000,
                                    _backend=harness.HostGraphBackend(2))
    assert got.tolist() == [want[i] for i in pick]


def test_read_do_and_write(tmp_path, capsys):
    src, out = tmp_path / "in.txt", tmp_path / "out.txt"
    src.write_text("[1, 0, 0, 0, 2, 0, 0, 0]\n[1, 2, 1,-2, 0, 0, 2, 1, -2, 0, 0, 0]\n[2, 1]\n")
    graph.read_do_and_write(str(src), str(out), 3, False, _backend=harness.HostGraphBackend(0))
    want = [bo.neighbourhood(r, 3, False) for r in ([1, 0, 0, 0, 2, 0, 0, 0], [1, 2, 1, -2, 0, 0, 2, 1, -2, 0, 0, 0], [2, 1])]
    assert [int(x) for x in out.read_text().split()] == want
    # README.txt:39-44: the reference prints the sorted presentation of every line
    assert capsys.readouterr().out.split("\n")[:3] == ["(a, b)", "(baB, abaB)", "(a, b)"]


def test_domain_errors():
    be = harness.HostGraphBackend(0)
    with pytest.raises(ValueError):
        graph.neighbourhood_sizes([[1, -1, 2, 0]], 2, True, _backend=be)  # not freely reduced
    with pytest.raises(ValueError):
        graph.neighbourhood_sizes([[0, 0, 2, 0]], 2, True, _backend=be)  # empty relator
    with pytest.raises(ValueError):
        graph.neighbourhood_sizes([[3, 2]], 2, True, _backend=be)  # letter outside the alphabet
    with pytest.raises(ValueError):
        graph.ac_graph_bfs(70, "classic", _backend=be)  # over the 56-letter planes of width 2
    # width 2 forced on a ball whose relators outgrow 56 letters: reported, never silently wrong
    long = [1, 2] * 12 + [2, 1] * 12
    with pytest.raises(OverflowError):
        graph.neighbourhood_sizes([long], 3, True, width_words=2, _backend=be)
    assert graph.neighbourhood_sizes([long], 2, True, _backend=be).tolist() == [bo.neighbourhood(long, 2, True)]


@pytest.mark.parametrize("width,maxlen", [(2, 27), (4, 59), (8, 123)])
def test_single_moves_equal_oracle_on_long_relators(width, maxlen):
    """g_move (multi-word plane shifts, junction cancellation of any depth, sort_) against Presentation::move of the
    oracle on random reduced words up to the plane capacity, incl. pairs built to cancel deeply."""
    from helpers import reduced_word

    rng = np.random.default_rng(width)
    cap = 32 * width - 8
    m = cap + 8
    rows, moves = [], []
    for k in range(3000):
        la, lb = int(rng.integers(1, maxlen + 1)), int(rng.integers(1, maxlen + 1))
        a, b = reduced_word(rng, la), reduced_word(rng, lb)
        kind = k % 4
        if kind == 1:  # b starts with the inverse of a's tail: deep cancellation in a.b
            t = int(rng.integers(1, min(la, lb) + 1))
            b = [-v for v in reversed(a[-t:])] + b[t:]
        elif kind == 2:  # b ends with the inverse of a's head: deep cancellation in b.a
            t = int(rng.integers(1, min(la, lb) + 1))
            b = b[:-t] + [-v for v in reversed(a[:t])]
        for w in (a, b):  # keep both freely reduced
            i = 0
            while i + 1 < len(w):
                if w[i] == -w[i + 1]:
                    del w[i:i + 2]
                    i = max(i - 1, 0)
                else:
                    i += 1
        if not a or not b:
            continue
        row = np.zeros(2 * m, dtype=np.int8)
        row[:len(a)] = a
        row[m:m + len(b)] = b
        rows.append(row)
        moves.append(int(rng.integers(0, 14)))
    rows = np.stack(rows)
    for classic in (True, False):
        mv = np.array(moves) % (14 if classic else 12)
        got, st = harness.graph_move(rows, mv, classic, width, allow_empty=True)
        n_ok = 0
        for j in range(len(rows)):
            want = bo.move(rows[j], int(mv[j]), classic)
            if want is None:  # empty relator (the oracle's row form cannot hold it) or longer than m
                continue
            child, size = want
            a_len, b_len = np.count_nonzero(child[:m]), np.count_nonzero(child[m:])
            if max(a_len, b_len) > cap:
                assert st[j] == 1, j
                continue
            assert st[j] == 0, (j, st[j])
            assert np.array_equal(got[j], child), (width, classic, j, int(mv[j]))
            n_ok += 1
        assert n_ok > 2000


def test_product_backend_needs_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(Exception):
        graph.ac_graph_bfs(4, "classic", device="cpu")
