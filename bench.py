#!/usr/bin/env python
"""bench.py -- AC-move transitions/s on B200 (BASELINE.json metric, config #2).

Workload ("step"): ONE lockstep ACEnv.step over a batch of 2^20 envs, max_relator_length = 36,
cyclical reduction, horizon 2000, auto-reset from the 1190 Miller-Schupp presentations.  States are
random AC walks (walk length U{0..64}) from the Miller-Schupp rows and AK(2), AK(3).

  value        device-timed transitions/s, states resident in HBM (CUDA events, max over ranks)
  e2e          the same step through the host-buffer API: actions come from pinned host memory,
               observations / reward / done / truncated go back to pinned host memory every step
  roofline     algorithmic bytes (4m+9 = 153 B per transition) / kernel time vs measured HBM peak
  cpu_baseline the CPU oracle (C restatement of the reference, oracle/) on the host cores

`--impl reference` times the CPU oracle port on the same workload instead (rank 0 only).
Multi-GPU: env batches shard with no communication (weak scaling: every rank steps its own 2^20 envs).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))

M = 36
N_ENV = 1 << 20
HORIZON = 2000
N_BATCHES = 4  # rotating resident batches: 4 x 75.5 MB of state > 126 MB L2
ALG_BYTES = 4 * M + 9  # read state 2m + action 1; write state 2m + lens 2 + reward 4 + done 1 + trunc 1
METRIC = "AC-move transitions/sec (lockstep ACEnv.step)"
UNIT = "transitions/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the env kernel, from the committed ncu capture
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_env_step_dram_traffic.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="acx", choices=["acx", "reference"])
    ap.add_argument("--envs", type=int, default=N_ENV)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python instead of one CUDA graph")
    ap.add_argument("--no-bfs", action="store_true", help="skip the hash-partitioned bfs leg (configs[3])")
    ap.add_argument("--no-other", action="store_true", help="skip configs[0], [2], [4] (one GPU only)")
    ap.add_argument("--bfs-budget", type=float, default=1e9)
    return ap.parse_args()


def config_dict(n_env, n_gpus):
    return {
        "workload": "BASELINE configs[1]: batched ACEnv.step, %d lockstep envs per GPU x random AC moves, "
                    "max_relator_length=36, cyclical reduction, horizon %d, auto-reset from the 1190 "
                    "Miller-Schupp presentations" % (n_env, HORIZON),
        "envs_per_gpu": n_env, "max_relator_length": M, "horizon": HORIZON,
        "l2": "rotating %d resident batches (%.0f MB of state per GPU) > 126 MB L2; no flush needed"
              % (N_BATCHES, N_BATCHES * n_env * 2 * M / 1e6),
        "sharding": "env batches sharded across %d GPU(s), no communication" % n_gpus,
        "reference_arm_envs": n_env,
    }


# ------------------------------------------------------------------------------------------------
def load_pool():
    """The 1190 Miller-Schupp presentations re-padded to m=36 (agents/environment.py:87-93) + AK(2), AK(3)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "dataset.npz"))
    widths, flat = z["widths"], z["flat"]
    off = np.concatenate([[0], np.cumsum(widths)])
    pool = np.zeros((len(widths) + 2, 2 * M), dtype=np.int8)
    for k in range(len(widths)):
        row = flat[off[k]:off[k + 1]]
        mm = len(row) // 2
        for h in range(2):
            r = row[h * mm:(h + 1) * mm]
            r = r[r != 0]
            pool[k, h * M:h * M + len(r)] = r
    pool[-2, :5] = [1, 1, -2, -2, -2]
    pool[-2, M:M + 6] = [1, 2, 1, -2, -1, -2]
    pool[-1, :7] = [1, 1, 1, -2, -2, -2, -2]
    pool[-1, M:M + 6] = [1, 2, 1, -2, -1, -2]
    return pool


def walk_states_gpu(pool_t, n, seed, dev):
    """n states: random AC walks of U{0..64} uniform moves from pool rows (generated with the engine)."""
    import torch
    from acx import batch

    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    start = torch.randint(0, pool_t.shape[0], (n,), device=dev, generator=g)
    states = pool_t[start].contiguous()
    walk = torch.randint(0, 65, (n,), device=dev, generator=g)
    for t in range(64):
        mv = torch.randint(0, 12, (n,), device=dev, generator=g, dtype=torch.int64).to(torch.uint8)
        nxt, _, st = batch.move_batch(states, mv, cyclical=True)
        live = (walk > t) & (st == 0)
        states = torch.where(live[:, None], nxt, states)
    return states.contiguous()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the GPU is under load (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 and len(r) >= 9] or [r for (_, r) in self.rows if len(r) >= 9]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[1]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({nm for r in rows for nm, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(float(r[3]) for r in rows)}


# ------------------------------------------------------------------------------------------------
def cpu_oracle_throughput(n_env, target_seconds, threads):
    """transitions/s of the CPU oracle (C restatement of the reference) on a bounded sample."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ac_oracle

    ac_oracle.build()
    pool = load_pool()
    rng = np.random.default_rng(0)
    n = min(n_env, 1 << 18)
    state = pool[rng.integers(0, len(pool), n)].copy()
    steps = np.zeros(n, dtype=np.int32)
    nxt = rng.integers(0, len(pool), n).astype(np.int32)
    for _ in range(16):  # spread the lengths a little, untimed
        ac_oracle.env_step_batch(state, steps, rng.integers(0, 12, n).astype(np.uint8), HORIZON, pool, nxt, threads)
    acts = [rng.integers(0, 12, n).astype(np.uint8) for _ in range(8)]
    done_steps, t0 = 0, time.perf_counter()
    while True:
        ac_oracle.env_step_batch(state, steps, acts[done_steps % 8], HORIZON, pool, nxt, threads)
        done_steps += 1
        dt = time.perf_counter() - t0
        if dt >= target_seconds:
            break
    return n * done_steps / dt, "%d ACEnv.step passes over %d envs (m=36) in %.1f s" % (done_steps, n, dt)


def reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores (rank 0 only).  The
    reference itself is pure Python; what is timed here is its C restatement (oracle/), all cores."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ac_oracle

    ac_oracle.build()
    cores = os.cpu_count() or 1
    pool = load_pool()
    rng = np.random.default_rng(0)
    n = args.envs  # the GPU arm's env count
    state = pool[rng.integers(0, len(pool), n)].copy()
    steps = np.zeros(n, dtype=np.int32)
    nxt = rng.integers(0, len(pool), n).astype(np.int32)
    acts = [rng.integers(0, 12, n).astype(np.uint8) for _ in range(8)]
    for k in range(max(args.warmup, 16)):
        ac_oracle.env_step_batch(state, steps, acts[k % 8], HORIZON, pool, nxt, cores)
    t0 = time.perf_counter()
    for k in range(args.steps):
        ac_oracle.env_step_batch(state, steps, acts[k % 8], HORIZON, pool, nxt, cores)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "i8", "data": "synthetic", "config": config_dict(args.envs, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "each step = one lockstep ACEnv.step pass over %d envs (m=36) with the C oracle "
                                   "(oracle/ac_oracle.c, pinned against the reference Python), %d pthreads" % (n, cores)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference itself is pure Python (10.7k ACMove/s/core, SURVEY.md section 6); this arm times "
                "its C restatement on all host cores, roughly 3 orders of magnitude faster than the reference",
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
def _time_steps(step, n_steps, warm0, dev, barrier, use_graph):
    """K launches timed with CUDA events on the launching stream; with use_graph the K launches are ONE CUDA graph
    (same kernels, same order, same buffers).  Returns (milliseconds, wall-clock start of the timed region)."""
    import torch

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    graph = None
    if use_graph:
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                for k in range(n_steps):
                    step(warm0 + k)
        torch.cuda.current_stream(dev).wait_stream(side)
    barrier()
    t_wall0 = time.time()
    e0.record()
    if graph is not None:
        graph.replay()
    else:
        for k in range(n_steps):
            step(warm0 + k)
    e1.record()
    barrier()
    return e0.elapsed_time(e1), t_wall0


def other_configs(dev):
    """configs[0], [2], [4] in the driver-run line (one GPU): same event discipline, a few seconds each"""
    import bench_configs as C

    out = {}
    g = C.cfg_greedy(dev)
    out["greedy_1190"] = {"config": g["config"], "wall_s": g["wall_s"], "solved": g["solved"], "expected_solved": 380,
                          "nodes_expanded_per_s": g["nodes_expanded_per_s"], "reference": g["reference"]}
    for key, kw in (("expand12_m128", dict(m=128, chunk=1 << 19, total=1 << 25)), ("expand12_m36", dict(m=36, chunk=1 << 20, total=1 << 26))):
        e = C.cfg_expand(dev, **kw)
        out[key] = {"config": e["config"].replace("1e8-state", "%d-state" % e["parents"]), "parents_per_s": e["parents_per_s"],
                    "transitions_per_s": e["transitions_per_s"], "frac": e["roofline"]["frac"], "alg_bytes_per_parent": e["alg_bytes_per_parent"]}
    out["ppo_rollout"] = C.cfg_ppo(dev)
    return out


def gpu_arm(args):
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # NCCL's INFO lines go to stdout on this image (NCCL_DEBUG_FILE is not honoured by the bundled library) and stdout
        # must stay the one JSON line: keep NCCL quiet and report the communicator's size on stderr instead (below)
        os.environ.setdefault("NCCL_DEBUG", "WARN")
    import torch
    import torch.distributed as dist

    from acx import batch, vecenv

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the AC-move engine has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        probe = torch.ones(1, device=dev)
        dist.all_reduce(probe)
        if rank == 0:
            sys.stderr.write("bench.py: NCCL communicator up: world_size %d, all_reduce(1) = %d, NCCL %s\n"
                             % (dist.get_world_size(), int(probe.item()), ".".join(map(str, torch.cuda.nccl.version()))))

    n = args.envs
    pool_np = load_pool()
    pool = torch.from_numpy(pool_np).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    bufs, acts, nxts = [], [], []
    for b in range(N_BATCHES):
        states = walk_states_gpu(pool, n, seed=100 * rank + b, dev=dev)
        # the timed step writes the 8-byte record only (reward, done, truncated, lengths); the clipped float reward and
        # the per-row status are optional outputs of the same kernel, used by the e2e leg below
        bufs.append(batch.EnvBuffers(states, HORIZON, pool=pool, want_clip=False, want_status=False))
        assert bufs[-1].certified, "bench states must be certified rows"
        bufs[-1].set_steps(torch.randint(0, HORIZON, (n,), device=dev, generator=g, dtype=torch.int32))
        acts.append(torch.randint(0, 12, (n,), device=dev, generator=g, dtype=torch.int64).to(torch.uint8))
        nxts.append(torch.randint(0, pool.shape[0], (n,), device=dev, generator=g, dtype=torch.int32))

    def make_step(chained):
        def step(k):
            b = k % N_BATCHES
            batch.env_step(bufs[b], acts[(k // N_BATCHES) % N_BATCHES], next_idx=nxts[b], chained=chained)
        return step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the actions of every step exist before the first launch and nothing else touches the env buffers between two
    # steps, so consecutive launches may be chained (per-CTA row dependency instead of the grid-wide one, acx.h)
    chained = os.environ.get("ACX_BENCH_CHAINED", "1") != "0"
    step = make_step(chained)
    # fraction of rows ONE step leaves unchanged (guard rejection, a conjugation that reduces back to the same word) on
    # the distribution the timed steps see: a scratch copy of batch 0 takes one step (61 % on uniform walks)
    scratch = batch.EnvBuffers(bufs[0].state.clone(), HORIZON, pool=pool, want_clip=False, want_status=False)
    scratch.set_steps(bufs[0].steps)
    before = scratch.state.clone()
    batch.env_step(scratch, acts[0], next_idx=nxts[0])
    torch.cuda.synchronize()
    noop = float((scratch.state == before).all(dim=1).float().mean().item())
    finished = float((scratch.done.bool() | scratch.truncated.bool()).float().mean().item())
    mean_total_len = float(scratch.lens.sum(dim=1).float().mean().item())
    del scratch, before
    for k in range(args.warmup):
        step(k)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    ms, t_wall0 = _time_steps(step, args.steps, args.warmup, dev, barrier, not args.no_graph)
    ms_unchained, _ = _time_steps(make_step(False), args.steps, args.warmup + args.steps, dev, barrier, not args.no_graph)
    # keep the same kernel running ~1.2 s more (untimed) so that nvidia-smi sees the clocks under this load
    t_hold = time.time()
    k = 0
    while time.time() - t_hold < 1.2:
        for _ in range(50):
            step(k)
            k += 1
        torch.cuda.synchronize()
    clocks = sampler.stop(t_wall0, time.time())

    # end to end through the host-buffer API (VecACEnv.step: numpy actions in, numpy results out)
    env = vecenv.VecACEnv(pool_np, horizon_length=HORIZON, num_envs=n, device=str(dev), track_actions=False,
                          start_index=np.random.default_rng(rank).integers(0, len(pool_np), n))
    env.reset()
    env.buf.set_state(bufs[1].state)
    host_actions = [np.random.default_rng(7 + i).integers(0, 12, n).astype(np.int64) for i in range(4)]
    e2e_steps = max(3, min(args.steps, 20))
    for i in range(3):
        env.step(host_actions[i % 4])
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        env.step(host_actions[i % 4])
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    tm = torch.tensor([ms, e2e_s * 1e3, ms_unchained], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms, e2e_ms, ms_unchained = float(tm[0]), float(tm[1]), float(tm[2])
    del bufs, acts, nxts, env
    torch.cuda.empty_cache()

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))

    bfs = None
    if not args.no_bfs:
        import bench_bfs

        if world == 1:
            bench_bfs.init_single_process_group(dev)
        bfs = bench_bfs.bfs_leg(dev, int(args.bfs_budget), peak_gbs=peak)
    others = None
    if world == 1 and not args.no_other and rank == 0:
        others = other_configs(dev)
    if rank != 0:
        if dist.is_initialized():
            dist.destroy_process_group()
        return

    value = world * n * args.steps / (ms * 1e-3)
    kern_s = ms * 1e-3 / args.steps  # one launch per step
    achieved = n * ALG_BYTES / kern_s / 1e9
    traffic, traffic_src = None, "no capture committed"
    try:
        with open(TRAFFIC_FILE) as f:
            tr = json.load(f)
        if n == tr.get("envs"):
            traffic, traffic_src = tr["dram_bytes_per_launch"], tr["source"]
    except (OSError, ValueError, KeyError):
        pass
    e2e_value = world * n * e2e_steps / (e2e_ms * 1e-3)
    d2h = n * 2 * M + 4 * n + 8 * n + 8
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "i8", "data": "synthetic", "config": config_dict(n, world),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n, "d2h_bytes_per_step": d2h,
                "api": "acx.vecenv.VecACEnv.step(numpy actions) -> numpy obs/reward/done/truncated/infos", "steps": e2e_steps,
                "bound": "pcie", "d2h_gbs_per_gpu": d2h / (e2e_ms * 1e-3 / e2e_steps) / 1e9,
                "note": "72 B of int8 observation per env per step cross PCIe (a host-side consumer cannot avoid them); the "
                        "kernel is 1-2 % of the step"},
        "gpu_launches": args.steps,
        "launch_mode": ("python loop" if args.no_graph else "one CUDA graph of the K launches") +
                       ("; consecutive steps chained by per-CTA row dependencies (ACX_ENV_CHAINED)" if chained else ""),
        "unchained": {"value": world * n * args.steps / (ms_unchained * 1e-3), "ms_per_step": ms_unchained / args.steps,
                      "frac": n * ALG_BYTES / (ms_unchained * 1e-3 / args.steps) / 1e9 / peak,
                      "note": "same launches ordered grid-after-grid (griddepcontrol.wait): what a step costs when its actions are "
                              "produced by the kernel launched just before it"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src,
                     "alg_bytes_per_launch": n * ALG_BYTES, "kernel": "k_env_tiles<MS=36,TR=128,S>",
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
                     "alg_bytes_per_transition": ALG_BYTES},
        "noop_row_fraction_per_step": noop, "finished_env_fraction_per_step": finished, "mean_total_length": mean_total_len,
    }
    if bfs is not None:
        line["bfs"] = bfs
    if others is not None:
        line["other_configs"] = others
    if world > 1:
        line["nccl"] = {"world_size": dist.get_world_size(), "version": ".".join(map(str, torch.cuda.nccl.version()))}
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        v, sample = cpu_oracle_throughput(n, 10.0 / max(1, cores) + 2.0, cores)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": sample + " (C oracle, %d pthreads; the reference itself is pure Python, "
                                                   "about 1e4 ACMove/s/core)" % cores}
    print(json.dumps(line))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        gpu_arm(a)
