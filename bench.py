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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="acx", choices=["acx", "reference"])
    ap.add_argument("--envs", type=int, default=N_ENV)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
