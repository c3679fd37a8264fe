"""GPU parity of acx_path_replay (ACMove along a time axis) against the oracle applied step by
step, and an end-to-end check: every greedy_search solution of the reference's 1190-presentation run replays to the
trivial presentation."""
import json
import os

import numpy as np
import pytest

from helpers import AK2, dataset_rows, random_rows, repad

pytestmark = pytest.mark.gpu


def _oracle_replay(oracle, roots, action_lists, cyc):
    P, rowb = roots.shape
    T = max([len(a) for a in action_lists], default=0)
    state = roots.copy()
    states = np.zeros((P, max(T, 1), rowb), dtype=np.int8)
    lens = np.zeros((P, max(T, 1), 2), dtype=np.int64)
    m = rowb // 2
    flens = np.stack([np.count_nonzero(roots[:, :m], axis=1), np.count_nonzero(roots[:, m:], axis=1)], axis=1)
    status = np.zeros(P, dtype=np.uint8)
    fail = np.full(P, -1, dtype=np.int64)
    for t in range(T):
        act = np.array([p for p in range(P) if len(action_lists[p]) > t and status[p] == 0], dtype=np.int64)
        if act.size == 0:
            break
        mv = np.array([action_lists[p][t] for p in act], dtype=np.uint8)
        out, ln, st = oracle.acmove_batch(state[act], mv, cyclical=cyc)
        for k, p in enumerate(act):
            if st[k]:
                status[p], fail[p] = st[k], t
            else:
                state[p] = out[k]
                states[p, t] = out[k]
                lens[p, t] = ln[k]
                flens[p] = ln[k]
    return states, lens, state, flens, status, fail


@pytest.mark.parametrize("m", [7, 18, 36, 64, 100, 128])
def test_replay_vs_oracle(cuda, oracle, golden_dir, m):
    from acx.replay import replay_paths

    rng = np.random.default_rng(1000 + m)
    for cyc in (True, False):
        rows = random_rows(rng, 400, m, kinds=(0, 1, 1, 2, 3, 4))
        if m >= 18:
            ds = [repad(r, m) for r in dataset_rows(golden_dir)[:: 13] if len(r) // 2 <= m]
            rows = np.concatenate([rows, np.stack(ds)])
        paths = [rng.integers(0, 12, size=int(rng.integers(0, 40))).tolist() for _ in range(len(rows))]
        paths[0] = []
        got = replay_paths(rows, paths, cyclical=cyc, device=str(cuda))
        states, lens, final, flens, status, fail = _oracle_replay(oracle, rows, paths, cyc)
        g_status, g_fail = got.status.cpu().numpy(), got.fail_step.cpu().numpy()
        assert (g_status == status).all()
        assert (g_fail == fail).all()
        assert (got.final.cpu().numpy() == final).all()
        ok = status == 0
        assert (got.final_lens.cpu().numpy()[ok] == flens[ok]).all()
        g_states, g_lens = got.states.cpu().numpy(), got.lens.cpu().numpy()
        for p in range(len(rows)):
            n = len(paths[p]) if status[p] == 0 else int(fail[p])
            assert (g_states[p, :n] == states[p, :n]).all(), (m, cyc, p)
            assert (g_lens[p, :n] == lens[p, :n]).all(), (m, cyc, p)
        assert status.any()  # the mix contains rows that fail somewhere along the way


def test_replay_reference_helpers(cuda, oracle):
    """replay_final_state / actions_to_path with the reference's signatures (agents/training.py:74-135)"""
    from acx.replay import actions_to_path, replay_final_state

    actions = [4, 0, 7, 9, 2, 11, 5, 1]
    state = AK2.copy()
    m = len(state) // 2
    want_path, want_detail = [], []
    for a in actions:
        state, ln = oracle.ACMove(a, state, m, None, True)
        want_path.append([a, int(sum(ln))])
        want_detail.append({"action": a, "r0": [int(x) for x in state[:m] if x], "r1": [int(x) for x in state[m:] if x],
                            "lengths": [int(ln[0]), int(ln[1])]})
    assert (replay_final_state(actions, AK2, str(cuda)) == state).all()
    assert actions_to_path(actions, AK2, str(cuda)) == (want_path, want_detail)
    with pytest.raises(AssertionError):  # r1 = r0^-1: move 3 (r1 -> r1 r0) empties r1
        replay_final_state([3], np.array([1, 2, 0, -2, -1, 0], dtype=np.int8), str(cuda))


def test_greedy_solutions_replay_to_trivial(cuda, golden_dir):
    """all 380 solution paths of greedy_search @ 10000 nodes (reference output) end in a length-2 presentation, and
    the length recorded at every step of every path is the replayed one"""
    from acx.replay import replay_paths

    with open(os.path.join(golden_dir, "greedy10k.json")) as f:
        g = json.load(f)
    rows = dataset_rows(golden_dir)
    idx = sorted(int(k) for k in g["paths"])
    assert len(idx) == 380
    roots36 = np.stack([repad(rows[i], 36) for i in idx])
    by_m = {}
    for i in idx:
        by_m.setdefault(len(rows[i]), []).append(i)
    for rowb, ids in by_m.items():
        roots = np.stack([rows[i] for i in ids])
        paths = [[int(a) for a, _ in g["paths"][str(i)][1:]] for i in ids]
        out = replay_paths(roots, paths, cyclical=False, device=str(cuda))
        assert not out.status.cpu().numpy().any()
        fl = out.final_lens.cpu().numpy()
        assert (fl.sum(axis=1) == 2).all()
        lens = out.lens.cpu().numpy().astype(np.int64).sum(axis=2)
        for k, i in enumerate(ids):
            want = [int(t) for _, t in g["paths"][str(i)][1:]]
            assert lens[k, : len(want)].tolist() == want
    assert roots36.shape[0] == 380
