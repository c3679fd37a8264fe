"""GPU: the drop-in `ac_solver` surface (reference names and signatures) against the vectors produced
by the reference's own tests (tests/golden/ref_test_tables.json) and its search / env goldens."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tables(golden_dir):
    with open(os.path.join(golden_dir, "ref_test_tables.json")) as f:
        return json.load(f)


def test_reference_tables_through_dropin(cuda, tables):
    from ac_solver.envs import ac_moves, utils

    for r in tables["simplify_relator"]:  # reference tests/test_ac_env.py:18-83
        out, n = utils.simplify_relator(np.array(r["relator"]), r["m"], r["cyclical"], r["padded"])
        assert n == r["len"] and np.array_equal(out, r["out"]), r
    for r in tables["is_array_valid_presentation"]:
        assert utils.is_array_valid_presentation(np.array(r["row"])) == r["out"], r
    for r in tables["is_presentation_trivial"]:
        assert utils.is_presentation_trivial(np.array(r["row"])) == r["out"], r
    for r in tables["simplify_presentation"]:
        out, lens = utils.simplify_presentation(np.array(r["row"]), r["m"], [0, 0], r["cyclical"])
        assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], r
    for name in ("concatenate_relators", "conjugate"):
        fn = getattr(ac_moves, name)
        for r in tables[name]:
            lengths = [0, 0]
            out, lens = fn(np.array(r["row"]), r["m"], r["i"], r["j"], r["sign"], lengths)
            assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], (name, r)
    for r in tables["ACMove"]:  # tests/test_ac_env.py:495-538
        row = np.array(r["row"])  # int64, as the reference test passes it
        if r["raises"]:
            with pytest.raises(AssertionError):
                ac_moves.ACMove(r["move"], row, r["m"], [0, 0], r["cyclical"])
        else:
            before = row.copy()
            out, lens = ac_moves.ACMove(r["move"], row, r["m"], [0, 0], r["cyclical"])
            assert np.array_equal(out, r["out"]) and list(lens) == r["lens"], r
            assert out.dtype == row.dtype and np.array_equal(row, before)


def test_helpers(cuda):
    from ac_solver.envs import utils

    t = utils.generate_trivial_states(3)
    assert t.shape == (8, 6) and all(utils.is_presentation_trivial(x) for x in t)
    p = utils.convert_relators_to_presentation([1, 2], [-2], 4)
    assert p.tolist() == [1, 2, 0, 0, -2, 0, 0, 0] and p.dtype == np.int8
    assert utils.change_max_relator_length_of_presentation(p, 2).tolist() == [1, 2, -2, 0]


def _cases(golden_dir, name, max_budget):
    with open(os.path.join(golden_dir, "search_golden.json")) as f:
        return [c for c in json.load(f)[name] if c["budget"] <= max_budget]


def test_greedy_search_golden(cuda, golden_dir, capsys):
    """incl. the exact AK(2) path of the reference's tests/search/test_gs.py and the budget-cut cases"""
    from ac_solver.search.greedy import greedy_search

    for c in _cases(golden_dir, "greedy", 10**6):
        ok, path = greedy_search(np.array(c["row"]), c["budget"], False, c["cyclical"])
        assert ok == c["solved"], c
        assert path == [tuple(p) for p in c["path"]], c
        out = capsys.readouterr().out
        if c["n_seen_at_exit"] is not None:
            assert f"explored nodes = {c['n_seen_at_exit']} " in out


def test_bfs_golden(cuda, golden_dir, capsys):
    from ac_solver.search.breadth_first import bfs

    for c in _cases(golden_dir, "bfs", 10**6):
        ok, path = bfs(np.array(c["row"]), c["budget"], False, c["cyclical"])
        assert ok == c["solved"], c
        assert path == (None if c["path"] is None else [tuple(p) for p in c["path"]]), c
        out = capsys.readouterr().out
        if c["n_seen_at_exit"] is not None:
            assert f"explored nodes = {c['n_seen_at_exit']} " in out


def test_greedy_all_1190_at_10k(cuda, golden_dir):
    """BASELINE config #1: greedy_search over all_presentations.txt at max_nodes_to_explore=10000:
    380 solved, every path identical to the reference's."""
    from acx import search
    from helpers import dataset_rows

    with open(os.path.join(golden_dir, "greedy10k.json")) as f:
        g = json.load(f)
    rows = dataset_rows(golden_dir)
    by_width = {}
    for k, r in enumerate(rows):
        by_width.setdefault(len(r), []).append(k)
    solved = np.zeros(len(rows), dtype=int)
    for width, idx in by_width.items():
        res, paths = search.search_batch(np.stack([rows[k] for k in idx]), 10000, False, True, str(cuda))
        assert (res[:, search.RES_ERROR] == 0).all()
        for j, k in enumerate(idx):
            solved[k] = res[j, search.RES_SOLVED]
            if solved[k]:
                assert paths[j] == [tuple(p) for p in g["paths"][str(k)]], k
            else:
                assert res[j, search.RES_SEEN] == g["n_seen_at_exit"][str(k)], k
    assert solved.tolist() == g["solved"] and solved.sum() == 380


def test_search_vs_oracle_random(cuda, oracle, golden_dir):
    from acx import search
    from helpers import dataset_rows

    rows = [r for r in dataset_rows(golden_dir) if len(r) == 36]  # the m=18 block
    rng = np.random.default_rng(3)
    for greedy in (True, False):
        for cyc in (False, True):
            for budget in (1, 13, 500, 3000):
                idx = [int(k) for k in rng.choice(len(rows), size=12, replace=False)]
                res, paths = search.search_batch(np.stack([rows[k] for k in idx]), budget, cyc, greedy, str(cuda))
                for j, k in enumerate(idx):
                    stats = {}
                    fn = oracle.greedy_search if greedy else oracle.bfs
                    ok, path = fn(rows[k], budget, False, cyc, stats=stats)
                    assert bool(res[j, search.RES_SOLVED]) == ok
                    assert paths[j] == path, (k, budget, cyc, greedy)
                    assert res[j, search.RES_SEEN] == stats["n_seen"]
                    assert res[j, search.RES_EXPANDED] == stats["n_expanded"]


def test_acenv_golden(cuda, golden_dir):
    """ACEnv.step / reset through the reference-shaped class, against trajectories of the reference"""
    from ac_solver import ACEnv, ACEnvConfig

    g = np.load(os.path.join(golden_dir, "env_golden.npz"))
    for e in (0, 5, 13, 30):
        env = ACEnv(ACEnvConfig(initial_state=g["init"][e].copy(), horizon_length=int(g["horizon"][e])))
        assert env.max_reward == int(g["horizon"][e]) * 36 * 2
        for t in range(120):
            obs, rew, done, trunc, info = env.step(int(g["actions"][t, e]))
            assert np.array_equal(obs, g["obs"][t, e]) and rew == g["reward"][t, e]
            assert done == bool(g["done"][t, e]) and trunc == bool(g["truncated"][t, e])
            assert env.lengths == g["lens"][t, e].tolist()
            assert ("actions" in info) == done
            if done or trunc:
                env.reset()
    with pytest.raises(ValueError):
        ACEnvConfig(initial_state=[1, 0, 0, 2, 0])
    with pytest.raises(ValueError):
        ACEnvConfig(initial_state=[0, 1, 2, 0])
    with pytest.raises(NotImplementedError):
        ACEnv(ACEnvConfig(use_supermoves=True))


def test_vec_env_host_api(cuda, oracle, golden_dir):
    """VecACEnv.step (numpy in / numpy out) with auto-reset, against the oracle"""
    from acx.vecenv import VecACEnv
    from helpers import dataset_rows, repad

    pool = np.stack([repad(r, 36) for r in dataset_rows(golden_dir)[:200]])
    n, horizon = 1000, 11
    env = VecACEnv(pool, horizon_length=horizon, num_envs=n, device=str(cuda))
    obs, _ = env.reset()
    state = pool[np.arange(n) % 200].copy()
    assert np.array_equal(obs, state)
    steps = np.zeros(n, dtype=np.int32)
    rng = np.random.default_rng(0)
    for t in range(30):
        act = rng.integers(0, 12, n)
        nxt = rng.integers(0, 200, n).astype(np.int32)
        env.set_next_states(nxt)
        o, r, d, tr, _ = env.step(act)
        rew, done, trunc, lens, st = oracle.env_step_batch(state, steps, act.astype(np.uint8), horizon, pool=pool, next_idx=nxt)
        assert np.array_equal(o, state) and np.array_equal(d, done.astype(bool)) and np.array_equal(tr, trunc.astype(bool))
        assert np.array_equal(r, np.clip(rew, -10, 1000).astype(np.float32))
        assert np.array_equal(env.lengths.cpu().numpy(), lens)
    env.envs[3].reset(options={"starting_state": pool[7]})
    assert np.array_equal(env.envs[3].state, pool[7]) and env.envs[3].count_steps == 0


def test_bfs_large_budget_level_engine(cuda, oracle, golden_dir, capsys):
    """budgets >= 2^16 go through the level-synchronous engine: same (solved, path), same exit message"""
    from ac_solver.search import breadth_first
    from helpers import AK3, dataset_rows, pad_row

    rows = [r for r in dataset_rows(golden_dir) if len(r) == 72]
    cases = [(pad_row(AK3[0], AK3[1], 36), 300000, False), (rows[len(rows) // 2], 80000, True), (rows[3], 70000, False)]
    for row, budget, cyc in cases:
        assert budget >= breadth_first.LEVEL_SYNCHRONOUS_FROM
        stats = {}
        ook, opath = oracle.bfs(row, budget, False, cyc, stats=stats)
        ok, path = breadth_first.bfs(row, budget, False, cyc)
        assert (ok, path) == (ook, opath)
        out = capsys.readouterr().out
        if not ok and stats["n_seen"] >= budget:
            assert f"explored nodes = {stats['n_seen']} has exceeded the limit {budget}" in out
