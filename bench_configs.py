#!/usr/bin/env python
"""bench_configs.py -- the other BASELINE.json configs (bench.py measures configs[1]).

    python bench_configs.py [greedy] [expand] [expand36] [ppo] [bfs] [replay] [graph [n]] [balls]

Prints one JSON line per config.  Same timing rules as bench.py (CUDA events, warm-up, footprint > L2)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402
from acx import batch, search  # noqa: E402

PEAK = 6538.3
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except OSError:
    pass


def dataset_rows():
    z = np.load(os.path.join(ROOT, "tests", "golden", "dataset.npz"))
    widths, flat = z["widths"], z["flat"]
    off = np.concatenate([[0], np.cumsum(widths)])
    return [flat[off[k]:off[k + 1]].copy() for k in range(len(widths))]


def cfg_greedy(dev):
    """configs[0]: greedy_search over the 1190 Miller-Schupp presentations, max_nodes_to_explore=10000."""
    rows = dataset_rows()
    by_width = {}
    for k, r in enumerate(rows):
        by_width.setdefault(len(r), []).append(k)
    groups = [np.stack([rows[k] for k in idx]) for idx in by_width.values()]
    for rep in range(2):  # first pass warms up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        solved = expanded = 0
        for g in groups:
            res, _ = search.search_batch(g, 10000, False, True, str(dev))
            solved += int(res[:, search.RES_SOLVED].sum())
            expanded += int(res[:, search.RES_EXPANDED].sum())
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    return {"config": "configs[0] greedy_search x 1190 presentations @ max_nodes_to_explore=10000", "solved": solved,
            "wall_s": dt, "nodes_expanded": expanded, "nodes_expanded_per_s": expanded / dt,
            "ac_moves_per_s": 12 * expanded / dt,
            "reference": "1356.8 s on one core, 380 solved (SURVEY.md section 6)", "includes": "H2D of roots, D2H of results/paths"}


def cfg_expand(dev, m=128, chunk=1 << 19, total=10 ** 8):
    """configs[4]: full 12-move neighbourhood expansion of a 1e8-state frontier, max_relator_length=128,
    streamed through HBM in chunks (children of one chunk: chunk*12*2m bytes)."""
    pool = torch.from_numpy(np.stack([np.concatenate([r[:len(r) // 2][r[:len(r) // 2] != 0],
                                                       np.zeros(m, np.int8)])[:m].tolist() +
                                      np.concatenate([r[len(r) // 2:][r[len(r) // 2:] != 0],
                                                      np.zeros(m, np.int8)])[:m].tolist()
                                      for r in dataset_rows()]).astype(np.int8)).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(2)
    nb = 3  # rotating parent chunks; the children buffers alone (chunk*12*2m) exceed L2 many times over
    parents = []
    for b in range(nb):
        states = pool[torch.randint(0, pool.shape[0], (chunk,), device=dev, generator=g)].contiguous()
        walk = torch.randint(0, 257, (chunk,), device=dev, generator=g)
        for t in range(256):
            mv = torch.randint(0, 12, (chunk,), device=dev, generator=g, dtype=torch.int64).to(torch.uint8)
            nxt, _, st = batch.move_batch(states, mv, cyclical=False)
            states = torch.where(((walk > t) & (st == 0))[:, None], nxt, states)
        parents.append(states.contiguous())
    n_chunks = (total + chunk - 1) // chunk
    for b in range(nb):
        batch.expand12(parents[b], cyclical=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    outs = [batch.expand12(parents[0], cyclical=False), batch.expand12(parents[1], cyclical=False)]
    import ctypes
    from acx import _lib
    L = _lib.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    torch.cuda.synchronize()
    e0.record()
    for k in range(n_chunks):
        ch, ln, st = outs[k % 2]
        L.acx_expand12(parents[k % nb].data_ptr(), ch.data_ptr(), ln.data_ptr(), st.data_ptr(), chunk, m, 0, stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    done = n_chunks * chunk
    alg = 26 * m + 24 + 12  # read 2m, write 12*(2m+2) (+12 status bytes)
    lens_mean = float(parents[0].ne(0).sum(dim=1).float().mean().item())
    return {"config": "configs[4] 12-move expansion of a 1e8-state frontier, max_relator_length=%d" % m,
            "parents": done, "chunk": chunk, "ms_total": ms, "parents_per_s": done / (ms * 1e-3),
            "transitions_per_s": 12 * done / (ms * 1e-3), "alg_bytes_per_parent": alg,
            "roofline": {"bound": "hbm", "achieved": done * alg / (ms * 1e-3) / 1e9, "peak": PEAK, "unit": "GB/s",
                         "frac": done * alg / (ms * 1e-3) / 1e9 / PEAK},
            "mean_total_length_of_parents": lens_mean}


def cfg_ppo(dev, n=65536, T=64):
    """configs[2]: PPO rollout with 65,536 GPU-resident envs.  The rollout of agents/training.py:229-241 as ONE CUDA
    graph of T steps: step t reads the observations from row t of the rollout buffer and writes row t+1 and the step
    record of row t (acx_env_io.state_out / step_out), so the env side of a step is exactly one kernel and no copy.
    env_only: pre-sampled actions, steps chained.  with_policy: the reference Agent's 2x256 tanh actor / critic
    (agents/ppo_agent.py:52-109) sample the actions on the same GPU inside the same graph."""
    from acx import vecenv

    pool = B.load_pool()
    env = vecenv.VecACEnv(pool, horizon_length=2000, num_envs=n, device=str(dev), track_actions=False)
    env.reset()
    buf = env.buf
    torch.manual_seed(0)
    actor = torch.nn.Sequential(torch.nn.Linear(72, 256), torch.nn.Tanh(), torch.nn.Linear(256, 256), torch.nn.Tanh(),
                                torch.nn.Linear(256, 12)).to(dev)
    critic = torch.nn.Sequential(torch.nn.Linear(72, 256), torch.nn.Tanh(), torch.nn.Linear(256, 256), torch.nn.Tanh(),
                                 torch.nn.Linear(256, 1)).to(dev)
    obs_buf = torch.zeros((T + 1, n, 72), dtype=torch.int8, device=dev)
    rec_buf = torch.zeros((T, n, 8), dtype=torch.uint8, device=dev)  # reward | done | truncated | lengths per step
    act_buf = torch.randint(0, 12, (T, n), device=dev, dtype=torch.int64).to(torch.uint8)
    val_buf = torch.zeros((T, n), dtype=torch.float32, device=dev)
    logp_buf = torch.zeros((T, n), dtype=torch.float32, device=dev)
    obs_buf[0].copy_(buf.state)

    def rollout(with_policy):
        for t in range(T):
            if with_policy:
                with torch.no_grad():
                    x = obs_buf[t].float()
                    logits = actor(x)
                    val_buf[t].copy_(critic(x).squeeze(-1))
                    logp = torch.log_softmax(logits, dim=-1)
                    # Categorical(logits).sample() as a Gumbel arg-max (one fused elementwise pass instead of multinomial)
                    a = (logp - torch.log(-torch.log(torch.rand_like(logp).clamp_min(1e-20)))).argmax(dim=-1)
                    logp_buf[t].copy_(logp.gather(1, a[:, None]).squeeze(-1))
                    act_buf[t].copy_(a)
            batch.env_step(buf, act_buf[t], next_idx=env.next_state_index, clip=env.clip, chained=not with_policy,
                           state_in=obs_buf[t], state_out=obs_buf[t + 1], rec_out=rec_buf[t])
        obs_buf[0].copy_(obs_buf[T])

    out = {}
    for with_policy in (False, True):
        key = "with_policy" if with_policy else "env_only"
        rollout(with_policy)  # warm-up, eager
        torch.cuda.synchronize()
        graph, mode = None, "eager"
        try:
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    rollout(with_policy)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph, mode = g, "one CUDA graph per rollout of %d steps" % T
        except Exception as ex:  # capture of the sampling ops is a torch matter: fall back to eager launches
            mode = "eager (graph capture failed: %s)" % type(ex).__name__
            torch.cuda.synchronize()
        reps = 5
        (graph.replay if graph is not None else (lambda: rollout(with_policy)))()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            if graph is not None:
                graph.replay()
            else:
                rollout(with_policy)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        out["env_steps_per_s_" + key] = reps * T * n / (ms * 1e-3)
        out["us_per_step_" + key] = 1e3 * ms / (reps * T)
        out["launch_mode_" + key] = mode
    assert int(buf.err_count.item()) == 0
    out.update({"config": "configs[2] PPO rollout, %d GPU-resident envs per B200, %d steps per rollout" % (n, T),
                "reference": "627 env-steps/s end to end with 4 envs (experiments/results/2026-02-23_03-11-29_ppo_rnd/results.json)"})
    return out


def cfg_replay(dev, n_paths=1 << 18, steps=64):
    """batched path replay (every intermediate state emitted), m=36."""
    from acx import _lib
    from acx.batch import _prep, _ptr

    rows = [r for r in dataset_rows() if len(r) == 72]
    rng = np.random.default_rng(5)
    roots = torch.from_numpy(np.stack([rows[i] for i in rng.integers(0, len(rows), size=n_paths)])).to(dev)
    acts = torch.from_numpy(rng.integers(0, 12, size=(n_paths, steps)).astype(np.uint8)).to(dev)
    states = torch.empty((n_paths, steps, 72), dtype=torch.int8, device=dev)
    lens = torch.empty((n_paths, steps, 2), dtype=torch.uint8, device=dev)
    final = torch.empty((n_paths, 72), dtype=torch.int8, device=dev)
    flen = torch.empty((n_paths, 2), dtype=torch.uint8, device=dev)
    fail = torch.empty((n_paths,), dtype=torch.int32, device=dev)
    status = torch.empty((n_paths,), dtype=torch.uint8, device=dev)
    stream = _prep(dev)
    L = _lib.lib()

    def run():
        _lib.check(L.acx_path_replay(_ptr(roots), _ptr(acts), 0, n_paths, steps, 36, 1, _ptr(states), _ptr(lens), _ptr(final),
                                     _ptr(flen), _ptr(fail), _ptr(status), stream), "acx_path_replay")

    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    alg = n_paths * steps * (72 + 2 + 1)  # every state and length pair written, one action read
    return {"config": "8(f) rank 2: path replay, %d paths x %d steps, max_relator_length=36, all intermediate states emitted" % (n_paths, steps),
            "ms": ms, "transitions_per_s": n_paths * steps / (ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": 6538.3, "unit": "GB/s",
                         "frac": alg / (ms * 1e-3) / 1e9 / 6538.3, "alg_bytes_per_step": 75},
            "failed_paths": int((status != 0).sum()),
            "reference": "one Python ACMove per step, about 1e4 per second per core (agents/training.py:74-135)"}


def cfg_graph(dev, n_max=14):
    """ac_bfs.cpp (classic moves) -- the whole graph of presentations of total size <= n."""
    from acx import graph

    out = []
    for moves in ("classic", "prime"):
        graph.ac_graph_bfs(6, moves, device=str(dev))  # warm-up (allocator, module load)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = graph.ac_graph_bfs(n_max, moves, device=str(dev), chunk=1 << 20)
        dt = time.perf_counter() - t0
        nm = 14 if moves == "classic" else 12
        out.append({"config": "8(f) rank 3: ac_bfs %s moves, n=%d (nodes, edges and filtrations on the host at exit)" % (moves, n_max),
                    "wall_s": dt, "nodes": len(r["sizes"]), "edges": len(r["edges"]), "levels": len(r["levels"]),
                    "nodes_per_s": len(r["sizes"]) / dt, "moves_per_s": nm * len(r["sizes"]) / dt,
                    "reference": "compiled ac_bfs.cpp (g++ -O2), classic: n=12 31.7 s, n=14 407 s (7199908 nodes, 20922303 edges) "
                                 "on one core of the build container"})
    return out


def cfg_balls(dev, radius=5):
    """neibourhoods.cpp -- radius-5 ball sizes around all 1190 Miller-Schupp presentations."""
    from acx import graph

    rows = [r.tolist() for r in dataset_rows()]
    out = []
    graph.neighbourhood_sizes(rows[:4], 3, True, device=str(dev))
    for classic in (True, False):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sizes = graph.neighbourhood_sizes(rows, radius, classic, device=str(dev), chunk=1 << 20)
        dt = time.perf_counter() - t0
        out.append({"config": "8(f) rank 3: neibourhoods %s moves, radius %d, 1190 Miller-Schupp presentations" % (
                        "classic" if classic else "prime", radius),
                    "wall_s": dt, "total_states": int(sizes.sum()), "states_per_s": float(sizes.sum()) / dt,
                    "min": int(sizes.min()), "max": int(sizes.max()), "checksum": int((sizes * np.arange(1, len(sizes) + 1)).sum()),
                    "reference": "compiled neibourhoods.cpp: about 1.5 s per presentation and move set on one core"})
    return out


if __name__ == "__main__":
    dev = torch.device("cuda:0")
    which = sys.argv[1:] or ["greedy", "expand", "ppo"]
    for w in which:
        if w == "greedy":
            print(json.dumps(cfg_greedy(dev)))
        elif w == "expand":
            print(json.dumps(cfg_expand(dev)))
        elif w == "expand36":
            print(json.dumps(cfg_expand(dev, m=36, chunk=1 << 20, total=1 << 26)))
        elif w == "ppo":
            print(json.dumps(cfg_ppo(dev)))
        elif w == "replay":
            print(json.dumps(cfg_replay(dev)))
        elif w == "graph":
            for line in cfg_graph(dev, int(os.environ.get("ACX_GRAPH_N", "14"))):
                print(json.dumps(line))
        elif w == "balls":
            for line in cfg_balls(dev):
                print(json.dumps(line))
        elif w == "bfs":
            import bench_bfs
            bench_bfs.init_single_process_group(dev)
            print(json.dumps(bench_bfs.bfs_leg(dev, 10 ** 7)))
