"""The CPU oracle against the golden vectors produced by RUNNING THE REFERENCE
(tests/golden/make_golden.py).  This is what pins the oracle; everything else is compared
with the oracle."""
import json
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def tables(golden_dir):
    with open(os.path.join(golden_dir, "ref_test_tables.json")) as f:
        return json.load(f)


def test_simplify_relator_table(oracle, tables):  # reference tests/test_ac_env.py:18-83
    for r in tables["simplify_relator"]:
        out, n = oracle.simplify_relator(np.array(r["relator"]), r["m"], r["cyclical"], r["padded"])
        assert n == r["len"] and np.array_equal(out, r["out"]), r


def test_predicate_tables(oracle, tables):  # tests/test_ac_env.py:86-124
    for r in tables["is_array_valid_presentation"]:
        assert oracle.is_array_valid_presentation(np.array(r["row"])) == r["out"], r
    for r in tables["is_presentation_trivial"]:
        assert oracle.is_presentation_trivial(np.array(r["row"])) == r["out"], r


def test_simplify_presentation_table(oracle, tables):  # tests/test_ac_env.py:140-181
    for r in tables["simplify_presentation"]:
        out, lens = oracle.simplify_presentation(np.array(r["row"]), r["m"], None, r["cyclical"])
        assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], r


def test_concat_conjugate_tables(oracle, tables):  # tests/test_ac_env.py:184-477
    for name in ("concatenate_relators", "conjugate"):
        fn = getattr(oracle, name)
        for r in tables[name]:
            out, lens = fn(np.array(r["row"]), r["m"], r["i"], r["j"], r["sign"], None)
            assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], (name, r)


def test_acmove_table(oracle, tables):  # tests/test_ac_env.py:495-538, all 12 moves, both flags
    for r in tables["ACMove"]:
        if r["raises"]:
            with pytest.raises(AssertionError):
                oracle.ACMove(r["move"], np.array(r["row"]), r["m"], None, r["cyclical"])
        else:
            out, lens = oracle.ACMove(r["move"], np.array(r["row"]), r["m"], None, r["cyclical"])
            assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], r


@pytest.mark.parametrize("m", [4, 7, 18, 36, 128])
def test_acmove_random_golden(oracle, golden_dir, m):
    g = np.load(os.path.join(golden_dir, "moves_golden.npz"))
    rows, mv, cyc, out, lens, raised = (g[f"{k}_{m}"] for k in ("rows", "moves", "cyc", "out", "lens", "raised"))
    for c in (0, 1):
        sel = cyc == c
        o, l, st = oracle.acmove_batch(rows[sel], mv[sel], cyclical=c)
        assert np.array_equal(o, out[sel])
        ok = raised[sel] == 0
        assert np.array_equal(l[ok], lens[sel][ok])
        assert np.array_equal(st != 0, raised[sel] != 0)


def test_env_golden(oracle, golden_dir):  # ACEnv.step trajectories produced by the reference
    g = np.load(os.path.join(golden_dir, "env_golden.npz"))
    init, horizon, actions = g["init"], g["horizon"], g["actions"]
    for hz in np.unique(horizon):
        sel = np.where(horizon == hz)[0]
        state = init[sel].copy()
        steps = np.zeros(len(sel), dtype=np.int32)
        idx = np.arange(len(sel), dtype=np.int32)
        for t in range(actions.shape[0]):
            rew, done, trunc, lens, st = oracle.env_step_batch(state, steps, actions[t, sel], int(hz))
            assert (st == 0).all()
            assert np.array_equal(state, g["obs"][t, sel])
            assert np.array_equal(rew, g["reward"][t, sel])
            assert np.array_equal(done, g["done"][t, sel])
            assert np.array_equal(trunc, g["truncated"][t, sel])
            assert np.array_equal(lens, g["lens"][t, sel])
            fin = (done | trunc).astype(bool)  # the generator resets finished envs to their start
            state[fin] = init[sel][fin]
            steps[fin] = 0
        del idx


def test_search_golden(oracle, golden_dir):
    with open(os.path.join(golden_dir, "search_golden.json")) as f:
        cases = json.load(f)
    for name, fn in (("greedy", oracle.greedy_search), ("bfs", oracle.bfs)):
        for c in cases[name]:
            if c["budget"] > 100000:
                continue  # the 1e6-budget cases are covered (slowly) by the GPU suite's goldens
            stats = {}
            ok, path = fn(np.array(c["row"], dtype=np.int8), c["budget"], False, c["cyclical"], stats=stats)
            assert ok == c["solved"], c
            exp = None if c["path"] is None else [tuple(p) for p in c["path"]]
            assert path == exp, c
            if c["n_seen_at_exit"] is not None:
                assert stats["n_seen"] == c["n_seen_at_exit"], c


def test_greedy10k_subsample(oracle, golden_dir):
    """BASELINE config #1 (380/1190 solved at 10K nodes): every 10th row here, all rows on the GPU box."""
    from helpers import dataset_rows

    with open(os.path.join(golden_dir, "greedy10k.json")) as f:
        g = json.load(f)
    assert sum(g["solved"]) == 380
    rows = dataset_rows(golden_dir)
    for k in range(0, len(rows), 10):
        ok, path = oracle.greedy_search(rows[k], 10000)
        assert int(ok) == g["solved"][k]
        if ok:
            assert path == [tuple(p) for p in g["paths"][str(k)]]
