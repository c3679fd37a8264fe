"""Generate the golden fixtures in tests/golden/ by RUNNING THE REFERENCE.

Build-container only (needs /root/reference; see oracle/ref_loader.py).  Every
file written here is an input/output record of the unmodified reference Python:
no reference source is copied.  Usage:

    python tests/golden/make_golden.py tables moves dataset search env
    python tests/golden/make_golden.py greedy10k        # ~5 min on 8 cores

Files:
  ref_test_tables.json  the (args -> result) rows of the reference's own
                        parametrized tests (tests/test_ac_env.py), with the result
                        RE-COMPUTED by the reference and checked against the
                        expectation the reference test states
  moves_golden.npz      seeded random (row, move, cyclical) -> (row', lengths, raised)
  dataset.npz           all_presentations.txt (1190 ragged rows) + solved lists +
                        the 533 stored greedy paths
  search_golden.json    greedy_search / bfs (solved, path, len(tree_nodes) at the
                        budget exit) on AK(2), AK(3), the MS cases of the reference's
                        search tests and random dataset rows
  env_golden.npz        ACEnv.step trajectories (obs, reward, done, truncated, lengths)
  env_info_golden.json  info["actions"] of the reference ACEnv at every done of those trajectories (target envinfo)
  greedy10k.json        BASELINE config #1: greedy_search on all 1190 rows at
                        max_nodes_to_explore=10000 (solved flag + path)
"""
import contextlib
import io
import json
import os
import re
import sys
import time
from ast import literal_eval

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_loader  # noqa: E402

AK2 = [1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0]
AK3_36 = [1, 1, 1, -2, -2, -2, -2] + [0] * 29 + [1, 2, 1, -2, -1, -2] + [0] * 30


def jsonable(x):
    if isinstance(x, np.ndarray):
        return x.tolist()
    if isinstance(x, (np.integer,)):
        return int(x)
    if isinstance(x, (np.bool_,)):
        return bool(x)
    if isinstance(x, (list, tuple)):
        return [jsonable(v) for v in x]
    return x


def read_rows(ref, name):
    with open(os.path.join(ref.root, "ac_solver/search/miller_schupp/data", name)) as f:
        return [literal_eval(line) for line in f if line.strip()]


def run_search(fn, row, budget, cyc):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ok, path = fn(np.array(row), budget, False, cyc)
    mt = re.search(r"explored nodes = (\d+)", buf.getvalue())
    return {
        "row": list(map(int, row)), "budget": int(budget), "cyclical": bool(cyc), "solved": bool(ok),
        "path": None if path is None else [[int(a), int(b)] for a, b in path],
        "n_seen_at_exit": int(mt.group(1)) if mt else None,
    }


# ---------------------------------------------------------------------------
def make_tables(ref):
    """Rows of the reference's parametrized tests, re-run through the reference."""
    import importlib

    sys.path.insert(0, ref.root)
    ref_loader._install_gymnasium_stand_in()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    try:
        tmod = importlib.import_module("tests.test_ac_env")
    finally:
        sys.path.remove(ref.root)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.") or k == "tests" or k.startswith("tests.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)

    def rows_of(name):
        fn = getattr(tmod, name)
        mark = [mk for mk in fn.pytestmark if mk.name == "parametrize"][0]
        return mark.args[1]

    out = {}
    t = []
    for rel, m, cyc, padded, exp, exp_n in rows_of("test_simplify_relator"):
        got, n = ref.simplify_relator(np.array(rel), m, cyc, padded)
        assert np.array_equal(got, exp) and n == exp_n
        t.append({"relator": jsonable(rel), "m": m, "cyclical": cyc, "padded": padded, "out": jsonable(got), "len": int(n)})
    out["simplify_relator"] = t
    t = []
    for pres, exp in rows_of("test_is_array_valid_presentation"):
        assert ref.is_array_valid_presentation(pres) == exp
        t.append({"row": jsonable(pres), "out": bool(exp)})
    out["is_array_valid_presentation"] = t
    t = []
    for pres, exp in rows_of("test_is_presentation_trivial"):
        assert ref.is_presentation_trivial(pres) == exp
        t.append({"row": jsonable(pres), "out": bool(exp)})
    out["is_presentation_trivial"] = t
    t = []
    for pres, m, lens, exp, exp_l in rows_of("test_simplify_presentation"):
        got, gl = ref.simplify_presentation(np.array(pres), m, list(lens))  # cyclical defaults to True
        assert np.array_equal(got, exp) and list(gl) == list(exp_l)
        t.append({"row": jsonable(pres), "m": m, "cyclical": True, "out": jsonable(got), "lens": jsonable(gl)})
        got, gl = ref.simplify_presentation(np.array(pres), m, list(lens), cyclical=False)
        t.append({"row": jsonable(pres), "m": m, "cyclical": False, "out": jsonable(got), "lens": jsonable(gl)})
    out["simplify_presentation"] = t
    for name, fn in (("concatenate_relators", ref.concatenate_relators), ("conjugate", ref.conjugate)):
        t = []
        for rels, nrel, i, j, sign, lens, exp, exp_l in rows_of("test_" + name):
            got, gl = fn(np.array(rels), nrel, i, j, sign, list(lens))
            assert np.array_equal(got, exp) and list(gl) == list(exp_l)
            t.append({"row": jsonable(rels), "m": nrel, "i": i, "j": j, "sign": sign, "out": jsonable(got), "lens": jsonable(gl)})
        out[name] = t
    # test_ACMove (tests/test_ac_env.py:495-538): all 12 moves on its state, both flags
    t = []
    state = np.array([1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0])
    for extra in ([1, 2, 0, -2, -1, 0], [2, 1, -2, 0, 0, 2, -1, 0, 0, 0], state.tolist()):
        s = np.array(extra)
        m = len(s) // 2
        for mv in range(12):
            for cyc in (True, False):
                try:
                    got, gl = ref.ACMove(mv, s, m, [int(np.count_nonzero(s[:m])), int(np.count_nonzero(s[m:]))], cyc)
                    t.append({"row": jsonable(s), "m": m, "move": mv, "cyclical": cyc, "out": jsonable(got), "lens": jsonable(gl), "raises": False})
                except AssertionError:
                    t.append({"row": jsonable(s), "m": m, "move": mv, "cyclical": cyc, "raises": True})
    out["ACMove"] = t
    with open(os.path.join(HERE, "ref_test_tables.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("tables:", {k: len(v) for k, v in out.items()})


def random_word(rng, n, reduced):
    w = []
    while len(w) < n:
        c = int(rng.choice([1, -1, 2, -2]))
        if reduced and w and w[-1] == -c:
            continue
        w.append(c)
    return w


def make_moves(ref):
    rng = np.random.default_rng(7)
    data = {}
    for m, count in ((4, 1500), (7, 1500), (18, 1500), (36, 4000), (128, 1200)):
        rows = np.zeros((count, 2 * m), dtype=np.int8)
        mv = rng.integers(0, 12, size=count).astype(np.uint8)
        cyc = rng.integers(0, 2, size=count).astype(np.uint8)
        out = np.zeros_like(rows)
        lens = np.zeros((count, 2), dtype=np.int32)
        raised = np.zeros(count, dtype=np.uint8)
        for k in range(count):
            kind = k % 3  # 0: random non-reduced, 1: reduced, 2: long (near the cap) reduced
            for h in range(2):
                n = int(rng.integers(1, m + 1)) if kind < 2 else int(rng.integers(max(1, m - 3), m + 1))
                rows[k, h * m : h * m + n] = random_word(rng, n, reduced=kind > 0)
            try:
                o, l = ref.ACMove(int(mv[k]), rows[k].copy(), m, [0, 0], cyclical=bool(cyc[k]))
                out[k], lens[k] = o, l
            except AssertionError:
                raised[k] = 1
                out[k] = rows[k]
        data.update({f"rows_{m}": rows, f"moves_{m}": mv, f"cyc_{m}": cyc, f"out_{m}": out,
                     f"lens_{m}": lens, f"raised_{m}": raised})
        print(f"moves m={m}: {count} cases, {int(raised.sum())} raise")
    np.savez_compressed(os.path.join(HERE, "moves_golden.npz"), **data)


def make_dataset(ref):
    rows = read_rows(ref, "all_presentations.txt")
    gs = read_rows(ref, "greedy_solved_presentations.txt")
    bs = read_rows(ref, "bfs_solved_presentations.txt")
    paths = read_rows(ref, "greedy_search_paths.txt")
    index = {tuple(r): k for k, r in enumerate(rows)}
    widths = np.array([len(r) for r in rows], dtype=np.int32)
    flat = np.concatenate([np.array(r, dtype=np.int8) for r in rows])
    p_off = np.cumsum([0] + [len(p) for p in paths]).astype(np.int32)
    p_flat = np.array([ab for p in paths for ab in p], dtype=np.int16).reshape(-1, 2)
    np.savez_compressed(
        os.path.join(HERE, "dataset.npz"), widths=widths, flat=flat,
        greedy_solved_idx=np.array([index[tuple(r)] for r in gs], dtype=np.int32),
        bfs_solved_idx=np.array([index[tuple(r)] for r in bs], dtype=np.int32),
        greedy_path_offsets=p_off, greedy_paths=p_flat,
    )
    print("dataset:", len(rows), "rows;", len(gs), "greedy-solved;", len(bs), "bfs-solved;", len(paths), "paths")


def ms_presentations(ref, n_lo, n_hi, w_lo, w_hi):
    import importlib

    sys.path.insert(0, ref.root)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    try:
        ms = importlib.import_module("ac_solver.search.miller_schupp.miller_schupp")
        out = []
        for n in range(n_lo, n_hi + 1):
            d = ms.generate_miller_schupp_presentations(n, w_hi)
            for w in range(w_lo, w_hi + 1):
                out += [list(map(int, p)) for p in d[w]]
        return out
    finally:
        sys.path.remove(ref.root)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)


def make_search(ref):
    rng = np.random.default_rng(11)
    rows = read_rows(ref, "all_presentations.txt")
    cases = {"greedy": [], "bfs": []}
    t0 = time.time()
    # tests/search/test_gs.py, test_bfs.py
    cases["greedy"].append(run_search(ref.greedy_search, AK2, int(1e6), False))
    cases["bfs"].append(run_search(ref.bfs, AK2, int(1e6), False))
    cases["bfs"].append(run_search(ref.bfs, AK2, 10, False))
    # tests/search/miller_schupp/test_miller_schupp.py:59-190
    for row in ms_presentations(ref, 1, 2, 1, 2):
        cases["greedy"].append(run_search(ref.greedy_search, row, int(1e6), False))
        cases["bfs"].append(run_search(ref.bfs, row, int(1e4), False))
    for row in ms_presentations(ref, 3, 4, 3, 4):
        cases["greedy"].append(run_search(ref.greedy_search, row, int(1e4), False))
    # budget-cut semantics on random rows, both flags
    for k in rng.choice(len(rows), size=30, replace=False):
        for budget in (10, 57, 300, 1000, 4000):
            cyc = bool(rng.integers(0, 2))
            cases["greedy"].append(run_search(ref.greedy_search, rows[int(k)], budget, cyc))
            cases["bfs"].append(run_search(ref.bfs, rows[int(k)], budget, cyc))
    # AK(3) at m=36 (BASELINE config #4 parity point)
    cases["bfs"].append(run_search(ref.bfs, AK3_36, 20000, False))
    cases["greedy"].append(run_search(ref.greedy_search, AK3_36, 20000, False))
    with open(os.path.join(HERE, "search_golden.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    print("search:", {k: len(v) for k, v in cases.items()}, f"{time.time() - t0:.0f}s")


def make_env(ref):
    rng = np.random.default_rng(13)
    rows = read_rows(ref, "all_presentations.txt")
    E, T = 48, 300
    m = 36
    init = np.zeros((E, 2 * m), dtype=np.int8)
    horizon = np.zeros(E, dtype=np.int32)
    actions = rng.integers(0, 12, size=(T, E)).astype(np.uint8)
    obs = np.zeros((T, E, 2 * m), dtype=np.int8)
    reward = np.zeros((T, E), dtype=np.int32)
    done = np.zeros((T, E), dtype=np.uint8)
    trunc = np.zeros((T, E), dtype=np.uint8)
    lens = np.zeros((T, E, 2), dtype=np.uint8)
    for e in range(E):
        base = np.array(rows[e % 12 if e < 24 else int(rng.integers(0, len(rows)))], dtype=np.int8)
        init[e] = ref.change_max_relator_length_of_presentation(base, m)
        horizon[e] = [7, 50, 1000][e % 3]
        env = ref.ACEnv(ref.ACEnvConfig(initial_state=init[e].copy(), horizon_length=int(horizon[e])))
        if e < 12:  # steer the easy envs along a cyclical=True greedy path so that `done` fires
            horizon[e] = 1000
            env = ref.ACEnv(ref.ACEnvConfig(initial_state=init[e].copy(), horizon_length=1000))
            with contextlib.redirect_stdout(io.StringIO()):
                ok, path = ref.greedy_search(init[e].copy(), 10000, False, True)
            assert ok
            acts = [a for a, _ in path[1:]]
            for start in (0, 100):
                actions[start : start + len(acts), e] = acts
        for t in range(T):
            o, r, d, tr, _ = env.step(int(actions[t, e]))
            obs[t, e], reward[t, e], done[t, e], trunc[t, e] = o, r, d, tr
            lens[t, e] = env.lengths
            if d or tr:  # reset to the initial state, as agents/training.py:351-352 does
                env.reset()
    np.savez_compressed(os.path.join(HERE, "env_golden.npz"), init=init, horizon=horizon, actions=actions,
                        obs=obs, reward=reward, done=done, truncated=trunc, lens=lens)
    print("env:", E, "envs x", T, "steps;", int(done.sum()), "done,", int(trunc.sum()), "truncated")


def make_envinfo(ref):
    """info["actions"] of the reference ACEnv (ac_env.py:92,110) at every `done` of the env_golden.npz trajectories:
    the same envs are re-run on the recorded actions with the reset-on-finish of agents/training.py:351-352."""
    g = np.load(os.path.join(HERE, "env_golden.npz"))
    init, horizon, actions = g["init"], g["horizon"], g["actions"]
    events = []
    for e in range(init.shape[0]):
        env = ref.ACEnv(ref.ACEnvConfig(initial_state=init[e].copy(), horizon_length=int(horizon[e])))
        for t in range(actions.shape[0]):
            o, r, d, tr, info = env.step(int(actions[t, e]))
            assert np.array_equal(o, g["obs"][t, e]) and bool(d) == bool(g["done"][t, e])
            if d:
                events.append({"t": t, "env": e, "actions": [int(a) for a in info["actions"]], "final_obs": jsonable(o)})
            else:
                assert info == {}
            if d or tr:
                env.reset()
    with open(os.path.join(HERE, "env_info_golden.json"), "w") as f:
        json.dump({"events": events}, f, separators=(",", ":"))
    print("envinfo:", len(events), "done events with action lists")


def _greedy_job(args):
    k, row = args
    ref = ref_loader.load()
    return k, run_search(ref.greedy_search, row, 10000, False)


def make_greedy10k(ref):
    import multiprocessing as mp

    rows = read_rows(ref, "all_presentations.txt")
    t0 = time.time()
    with mp.Pool(os.cpu_count()) as pool:
        res = dict(pool.imap_unordered(_greedy_job, list(enumerate(rows)), chunksize=4))
    out = {"budget": 10000, "solved": [], "paths": {}, "n_seen_at_exit": {}}
    for k in range(len(rows)):
        r = res[k]
        out["solved"].append(int(r["solved"]))
        if r["solved"]:
            out["paths"][str(k)] = r["path"]
        else:
            out["n_seen_at_exit"][str(k)] = r["n_seen_at_exit"]
    with open(os.path.join(HERE, "greedy10k.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("greedy10k:", sum(out["solved"]), "solved of", len(rows), f"{time.time() - t0:.0f}s")


if __name__ == "__main__":
    ref = ref_loader.load()
    for what in sys.argv[1:]:
        globals()["make_" + what](ref)
