"""Env-vector layer on the GPU: action log / infos on done (ac_env.py:92,110 through gymnasium's final_info layout,
agents/training.py:273-280), launch chaining of the certified tile kernel, and the reference rollout loop's
done / truncated handling (training.py:258-352) driven over get_env / VecACEnv."""
import json
import os
import random
import types

import numpy as np
import pytest

from helpers import dataset_rows, repad, walk_states

pytestmark = pytest.mark.gpu


def _t(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def test_vecenv_infos_match_reference_action_lists(cuda, golden_dir):
    """the reference-recorded trajectories of env_golden.npz through VecACEnv.step: at every `done` the action list
    in infos equals the reference ACEnv's info["actions"] (tests/golden/env_info_golden.json), final_observation is
    the pre-reset state, and a truncated env carries an empty final_info"""
    from acx.vecenv import VecACEnv

    g = np.load(os.path.join(golden_dir, "env_golden.npz"))
    with open(os.path.join(golden_dir, "env_info_golden.json")) as f:
        events = {(ev["t"], ev["env"]): ev for ev in json.load(f)["events"]}
    init, horizon, actions = g["init"], g["horizon"], g["actions"]
    seen = 0
    for hz in np.unique(horizon):
        sel = np.where(horizon == hz)[0]
        envs = VecACEnv(init[sel], horizon_length=int(hz), clip_rewards=False, device=str(cuda))
        assert envs.track_actions
        envs.reset()
        for t in range(actions.shape[0]):
            obs, rew, done, trunc, infos = envs.step(actions[t, sel].astype(np.int64))
            assert np.array_equal(done, g["done"][t, sel].astype(bool))
            assert np.array_equal(trunc, g["truncated"][t, sel].astype(bool))
            assert np.array_equal(rew, g["reward"][t, sel].astype(np.float32))
            fin = done | trunc
            exp = g["obs"][t, sel].copy()
            exp[fin] = init[sel][fin]  # re-seated on its own initial state (SyncVectorEnv auto-reset)
            assert np.array_equal(obs, exp)
            if not fin.any():
                assert infos == {}
                continue
            assert "final_info" in infos and "final_observation" in infos
            for j, e in enumerate(sel):
                if not fin[j]:
                    assert infos["final_info"][j] is None and not infos["_final_info"][j]
                    continue
                assert infos["_final_info"][j] and infos["_final_observation"][j]
                assert np.array_equal(infos["final_observation"][j], g["obs"][t, e])
                if done[j]:
                    ev = events[(t, int(e))]
                    assert infos["final_info"][j]["actions"] == ev["actions"]
                    assert list(infos["actions"][j]) == ev["actions"] and infos["_actions"][j]
                    seen += 1
                else:
                    assert infos["final_info"][j] == {}
    assert seen == len(events) > 0


def test_env_handle_surface(cuda, golden_dir):
    """envs.envs[i]: what agents/training.py:352,362-363 touches"""
    from acx.vecenv import VecACEnv

    roots = dataset_rows(golden_dir)
    pool = np.stack([repad(r, 36) for r in roots[:8]])
    envs = VecACEnv(pool, horizon_length=50, device=str(cuda))
    envs.reset()
    h = envs.envs[3]
    assert h.unwrapped is h and h.max_reward == 50 * 36 * 2 and h.horizon_length == 50
    acts = [4, 7, 0, 11]
    for a in acts:
        envs.step(np.full(8, a))
    assert h.actions == acts and h.count_steps == 4
    st, _ = h.reset(options={"starting_state": pool[5]})
    assert np.array_equal(st, pool[5]) and h.count_steps == 0 and h.actions == []
    assert h.lengths == [int(np.count_nonzero(pool[5][:36])), int(np.count_nonzero(pool[5][36:]))]
    assert envs.single_action_space.shape == () and envs.single_observation_space.shape == (72,)


def test_vecenv_step_raises_like_the_reference(cuda):
    """r_1 == r_0^-1: the concatenation empties a relator and the reference's ACMove asserts (utils.py:264-266)"""
    from acx.vecenv import VecACEnv

    row = np.zeros(72, dtype=np.int8)
    row[:2] = [1, 2]
    row[36:38] = [-2, -1]
    envs = VecACEnv(np.stack([row, row]), horizon_length=20, device=str(cuda))
    envs.reset()
    with pytest.raises(AssertionError):
        envs.step(np.array([3, 4]))
    envs.reset()
    import torch

    envs.step_device(torch.tensor([3, 4], dtype=torch.uint8, device=cuda))
    with pytest.raises(AssertionError):
        envs.check_errors()


@pytest.mark.parametrize("n", [128 * 40, 128 * 1184 + 256, 70001])
def test_env_step_chained_graph(cuda, oracle, golden_dir, n):
    """T chained steps (per-CTA dependency instead of the grid-wide one) captured in ONE CUDA graph and replayed
    twice give the trajectory of T (then 2T) ordinary steps, and match the oracle"""
    import torch
    from acx import batch

    m, horizon, T = 36, 11, 24
    roots = dataset_rows(golden_dir)
    pool = np.stack([repad(r, m) for r in roots[:200]])
    states = walk_states(oracle, roots, m, 4096, 30, seed=3, cyclical=True)
    rng = np.random.default_rng(n)
    state = states[rng.integers(0, len(states), n)].copy()
    steps = np.zeros(n, dtype=np.int32)
    acts = [rng.integers(0, 12, n).astype(np.uint8) for _ in range(T)]
    nxt = rng.integers(0, len(pool), n).astype(np.int32)
    a = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(pool, cuda), want_clip=False, want_status=False)
    b = batch.EnvBuffers(_t(state, cuda), horizon=horizon, pool=_t(pool, cuda))
    assert a.certified
    d_acts = [_t(x, cuda) for x in acts]
    d_nxt = _t(nxt, cuda)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream(cuda)
    side.wait_stream(torch.cuda.current_stream(cuda))
    with torch.cuda.stream(side):
        batch.env_step(a, d_acts[0], next_idx=d_nxt)  # warm-up outside the capture (function attributes)
        a.set_state(_t(state, cuda), torch.zeros(n, dtype=torch.int32, device=cuda))
        with torch.cuda.graph(graph, stream=side):
            for t in range(T):
                batch.env_step(a, d_acts[t], next_idx=d_nxt, chained=True)
    torch.cuda.current_stream(cuda).wait_stream(side)
    for rep in range(2):
        graph.replay()
        for t in range(T):
            batch.env_step(b, d_acts[t], next_idx=d_nxt)
            rew, done, trunc, lens, st = oracle.env_step_batch(state, steps, acts[t], horizon, pool=pool, next_idx=nxt, threads=8)
        torch.cuda.synchronize()
        assert np.array_equal(b.state.cpu().numpy(), state)
        assert np.array_equal(a.state.cpu().numpy(), state)
        assert np.array_equal(a.meta.cpu().numpy(), b.meta.cpu().numpy())
        assert np.array_equal(a.step_out.cpu().numpy(), b.step_out.cpu().numpy())
        assert np.array_equal(a.steps.cpu().numpy(), steps)
        assert np.array_equal(a.reward.cpu().numpy(), rew)
        assert int(a.err_count.item()) == 0


def test_rollout_loop_over_get_env(cuda, golden_dir):
    """the reference rollout's bookkeeping (agents/training.py:229-352: episodic counters, ACMoves_hist from the
    action lists, curriculum reset through envs.envs[i].reset) driven over get_env with scripted actions: the envs that
    follow their stored greedy path are recorded as solved with exactly that path"""
    import torch
    from ac_solver.agents.environment import get_env

    z = np.load(os.path.join(golden_dir, "dataset.npz"))
    args = types.SimpleNamespace(fixed_init_state=False, states_type="all", num_envs=16, horizon_length=40,
                                 clip_rewards=True, min_rew=-10, max_rew=1000, norm_rewards=False, use_supermoves=False,
                                 device=str(cuda), repeat_solved_prob=0.25, max_relator_length=36)
    envs, initial_states, curr_states, success_record, ACMoves_hist, states_processed = get_env(args)
    n = args.num_envs
    # scripted policy: env i replays the cyclical greedy path of its current presentation when the oracle has one
    import ac_oracle

    paths = {}

    def plan(idx):
        if idx not in paths:
            ok, path = ac_oracle.greedy_search(np.asarray(initial_states[idx], dtype=np.int8), 2000,
                                               cyclically_reduce_after_moves=True)
            paths[idx] = [a for a, _ in path[1:]] if ok else None
        return paths[idx]

    rng = random.Random(0)
    next_obs = torch.Tensor(envs.reset()[0]).to(cuda)
    episodic_return, episodic_length = np.zeros(n), np.zeros(n, dtype=np.int64)
    pos = [0] * n
    round1_complete = False
    solved_with = {}
    for step in range(160):
        action = np.zeros(n, dtype=np.int64)
        for i in range(n):
            p = plan(curr_states[i])
            action[i] = p[pos[i]] if (p is not None and pos[i] < len(p)) else rng.randrange(12)
            pos[i] += 1
        next_obs, reward, done, truncated, infos = envs.step(action)
        assert torch.tensor(reward).to(cuda).view(-1).shape == (n,)
        episodic_return = episodic_return + reward
        episodic_length = episodic_length + 1
        record = np.array([bool(done[i] or truncated[i]) for i in range(n)])
        if record.any():
            for i, el in enumerate(record):
                if done[i]:
                    if curr_states[i] in success_record["unsolved"]:
                        success_record["unsolved"].remove(curr_states[i])
                        success_record["solved"].add(curr_states[i])
                    if "final_info" in infos:
                        action_list = infos["final_info"][i]["actions"]
                    else:
                        action_list = list(infos["actions"][i])
                    assert len(action_list) == episodic_length[i]
                    if curr_states[i] not in ACMoves_hist or len(action_list) < len(ACMoves_hist[curr_states[i]]):
                        ACMoves_hist[curr_states[i]] = action_list
                    solved_with[curr_states[i]] = action_list
                if el:
                    episodic_return[i], episodic_length[i] = 0, 0
                    round1_complete = round1_complete or (max(states_processed) == len(initial_states) - 1)
                    if not round1_complete:
                        curr_states[i] = max(states_processed) + 1
                    elif len(success_record["solved"]) == 0 or (success_record["unsolved"] and rng.uniform(0, 1) > args.repeat_solved_prob):
                        curr_states[i] = rng.choice(list(success_record["unsolved"]))
                    else:
                        curr_states[i] = rng.choice(list(success_record["solved"]))
                    states_processed.add(curr_states[i])
                    next_obs[i] = initial_states[curr_states[i]]
                    envs.envs[i].reset(options={"starting_state": next_obs[i]})
                    pos[i] = 0
        next_obs = torch.Tensor(next_obs).to(cuda)
    base_env = envs.envs[0].unwrapped if hasattr(envs.envs[0], "unwrapped") else envs.envs[0]
    assert base_env.max_reward == 40 * 36 * 2
    assert solved_with, "no scripted env reached the trivial presentation"
    for idx, lst in solved_with.items():
        assert lst == plan(idx)  # the action list handed back is the path that was played
        # ... and it does trivialise the presentation (agents/training.py:297-303 replays it the same way)
        row = np.asarray(initial_states[idx], dtype=np.int8)[None, :].copy()
        for a in lst:
            row, lens, st = ac_oracle.acmove_batch(row, np.array([a], dtype=np.uint8), cyclical=1, threads=1)
        assert int(lens.sum()) == 2
    del z
