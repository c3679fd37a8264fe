"""Pin the C oracle against the reference Python, run in the build container.

TEST INFRASTRUCTURE ONLY.  Usage: python oracle/validate_vs_reference.py [--quick]
Compares, on seeded random inputs (reduced walks AND non-reduced rows, several
max_relator_length values, both `cyclical` settings):
  ACMove / concatenate_relators / conjugate / simplify_relator /
  simplify_presentation / predicates, ACEnv.step trajectories,
  greedy_search and bfs (solved flag + exact path) on dataset rows.
Exits non-zero on the first mismatch.  The last run's summary is committed as
oracle/VALIDATION.txt.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ac_oracle as orc  # noqa: E402
import ref_loader  # noqa: E402


def random_word(rng, n, alphabet, reduced):
    w = []
    while len(w) < n:
        c = int(rng.choice(alphabet))
        if reduced and w and w[-1] == -c:
            continue
        w.append(c)
    return w


def random_row(rng, m, alphabet=(1, -1, 2, -2), reduced=False):
    row = np.zeros(2 * m, dtype=np.int8)
    for h in range(2):
        n = int(rng.integers(1, m + 1))
        row[h * m : h * m + n] = random_word(rng, n, alphabet, reduced)
    return row


def check_moves(ref, rng, n_cases, log):
    t0 = time.time()
    raised = 0
    for case in range(n_cases):
        m = int(rng.choice([2, 3, 4, 7, 12, 18, 36, 40, 128]))
        reduced = bool(rng.integers(0, 2))
        row = random_row(rng, m, reduced=reduced)
        mv = int(rng.integers(0, 12))
        cyc = bool(rng.integers(0, 2))
        lens = [int(np.count_nonzero(row[:m])), int(np.count_nonzero(row[m:]))]
        try:
            exp, exp_l = ref.ACMove(mv, row.copy(), m, list(lens), cyclical=cyc)
            exp_raise = False
        except AssertionError:
            exp_raise = True
        got, got_l, st = orc.acmove_batch(row[None, :], [mv], cyclical=cyc)
        if exp_raise:
            raised += 1
            assert st[0] == 1, (case, row, mv, cyc, st)
        else:
            assert st[0] == 0, (case, row, mv, cyc, st)
            assert np.array_equal(got[0], exp) and list(got_l[0]) == [int(x) for x in exp_l], (
                case, row.tolist(), mv, m, cyc, got[0].tolist(), exp.tolist())
    log(f"ACMove: {n_cases} random (row, move, m, cyclical) cases identical "
        f"({raised} where the reference raises AssertionError -> status 1) [{time.time() - t0:.1f}s]")


def check_walks(ref, rng, n_walks, steps, log):
    """Random walks from dataset rows: the regime every benchmark row lives in."""
    rows = load_dataset(ref)
    n = 0
    for w in range(n_walks):
        base = rows[int(rng.integers(0, len(rows)))]
        m0 = len(base) // 2
        m = int(rng.choice([m0, 36, 48]))
        if m < m0:
            m = m0
        s = ref.change_max_relator_length_of_presentation(np.array(base, dtype=np.int8), m)
        cyc = bool(w % 2)
        for _ in range(steps):
            mv = int(rng.integers(0, 12))
            exp, exp_l = ref.ACMove(mv, s, m, [0, 0], cyclical=cyc)
            got, got_l = orc.ACMove(mv, s, m, None, cyclical=cyc)
            assert np.array_equal(got, exp) and got_l == [int(x) for x in exp_l], (s.tolist(), mv, m, cyc)
            s = exp
            n += 1
    log(f"ACMove walks: {n} steps over {n_walks} walks from all_presentations.txt rows identical")


def check_words(ref, rng, n_cases, log):
    for case in range(n_cases):
        m = int(rng.choice([3, 5, 8, 20]))
        alphabet = (1, -1, 2, -2, 3, -3, 4, -4)
        n = int(rng.integers(0, m + 1))
        w = np.zeros(m, dtype=np.int64)
        w[:n] = random_word(rng, n, alphabet, reduced=False)
        for cyc in (False, True):
            for padded in (False, True):
                exp, exp_n = ref.simplify_relator(w.copy(), m, cyclical=cyc, padded=padded)
                got, got_n = orc.simplify_relator(w.copy(), m, cyclical=cyc, padded=padded)
                assert exp_n == got_n and np.array_equal(exp, got), (w, cyc, padded, exp, got)
        row = random_row(rng, m, alphabet)
        lens = [int(np.count_nonzero(row[:m])), int(np.count_nonzero(row[m:]))]
        for i in (0, 1):
            for sign in (1, -1):
                exp, _ = ref.concatenate_relators(row.copy(), m, i, 1 - i, sign, list(lens))
                got, got_l = orc.concatenate_relators(row.copy(), m, i, 1 - i, sign, None)
                assert np.array_equal(exp, got), (row, i, sign, exp, got)
                for j in (1, 2):
                    exp, exp_l = ref.conjugate(row.copy(), m, i, j, sign, list(lens))
                    got, got_l = orc.conjugate(row.copy(), m, i, j, sign, None)
                    assert np.array_equal(exp, got) and [int(x) for x in exp_l] == got_l, (row, i, j, sign, exp, got)
        for cyc in (False, True):
            exp, exp_l = ref.simplify_presentation(row.copy(), m, list(lens), cyclical=cyc)
            got, got_l = orc.simplify_presentation(row.copy(), m, None, cyclical=cyc)
            assert np.array_equal(exp, got) and [int(x) for x in exp_l] == got_l
        # predicates on arbitrary (possibly invalid) rows over {-2..2}
        junk = rng.integers(-2, 3, size=2 * int(rng.integers(1, 5))).astype(np.int8)
        assert bool(ref.is_array_valid_presentation(junk)) == orc.is_array_valid_presentation(junk), junk
        assert bool(ref.is_presentation_trivial(junk)) == orc.is_presentation_trivial(junk), junk
    log(f"simplify_relator / concatenate_relators / conjugate / simplify_presentation / predicates: "
        f"{n_cases} random cases (alphabet +-1..+-4, non-reduced) identical")


def check_env(ref, rng, n_envs, steps, log):
    rows = load_dataset(ref)
    n_done = 0
    for e in range(n_envs):
        base = np.array(rows[int(rng.integers(0, len(rows)))], dtype=np.int8)
        if e % 3 == 0:  # easy rows so that `done` fires
            base = np.array(rows[e % 8], dtype=np.int8)
        m = len(base) // 2
        horizon = int(rng.choice([5, 17, 1000]))
        env = ref.ACEnv(ref.ACEnvConfig(initial_state=base.copy(), horizon_length=horizon))
        s = base.copy()[None, :]
        st = np.zeros(1, dtype=np.int32)
        for _ in range(steps):
            a = int(rng.integers(0, 12))
            obs, rew, done, trunc, _ = env.step(a)
            r, d, t, l, status = orc.env_step_batch(s, st, [a], horizon)
            assert np.array_equal(obs, s[0]) and int(rew) == int(r[0]) and bool(done) == bool(d[0]) \
                and bool(trunc) == bool(t[0]) and [int(x) for x in env.lengths] == list(l[0]), (base, a)
            n_done += bool(done)
            if done or trunc:
                env.reset()
                s[0] = base
                st[0] = 0
    log(f"ACEnv.step: {n_envs} envs x {steps} steps identical (obs, reward, done, truncated, lengths); "
        f"{n_done} done events")


def load_dataset(ref):
    from ast import literal_eval

    path = os.path.join(ref.root, "ac_solver/search/miller_schupp/data/all_presentations.txt")
    with open(path) as f:
        return [literal_eval(line) for line in f if line.strip()]


def check_search(ref, rng, n_rows, budgets, log):
    import contextlib
    import io

    rows = load_dataset(ref)
    ak2 = [1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0]
    picks = [ak2] + [rows[int(i)] for i in rng.choice(len(rows), size=n_rows, replace=False)]
    n = 0
    t0 = time.time()
    for row in picks:
        for budget in budgets:
            for cyc in (False, True):
                with contextlib.redirect_stdout(io.StringIO()):
                    e_ok, e_path = ref.greedy_search(np.array(row), budget, False, cyc)
                    b_ok, b_path = ref.bfs(np.array(row), budget, False, cyc)
                g_ok, g_path = orc.greedy_search(row, budget, False, cyc)
                assert g_ok == e_ok and g_path == [(int(a), int(b)) for a, b in e_path], (row, budget, cyc)
                o_ok, o_path = orc.bfs(row, budget, False, cyc)
                assert o_ok == b_ok, (row, budget, cyc)
                if b_ok:
                    assert o_path == [(int(a), int(b)) for a, b in b_path], (row, budget, cyc)
                else:
                    assert o_path is None and b_path is None
                n += 2
    log(f"greedy_search + bfs: {n} (row, budget, cyclical) searches identical in (solved, exact path) "
        f"[budgets {budgets}; {time.time() - t0:.1f}s]")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    ref = ref_loader.load()
    orc.build()
    lines = []

    def log(msg):
        print(msg, flush=True)
        lines.append(msg)

    rng = np.random.default_rng(20261104)
    q = args.quick
    check_moves(ref, rng, 3000 if q else 60000, log)
    check_walks(ref, rng, 20 if q else 200, 100 if q else 300, log)
    check_words(ref, rng, 200 if q else 3000, log)
    check_env(ref, rng, 10 if q else 60, 100 if q else 400, log)
    check_search(ref, rng, 3 if q else 24, [10, 57, 300] if q else [10, 57, 300, 1000, 4000], log)
    log("ALL IDENTICAL")
    if not q:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "VALIDATION.txt"), "w") as f:
            f.write("oracle/validate_vs_reference.py, full run against /root/reference (numpy %s)\n" % np.__version__)
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
