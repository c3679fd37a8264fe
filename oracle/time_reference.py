"""Times the UNMODIFIED Python reference (imported from /root/reference; build container only) on the hot path:
ACMove / ACEnv.step on one core and over multiprocessing.Pool (the reference's own scaling mechanism,
scripts/parallel_search.py:202-203), and bfs from AK(3).  TEST INFRASTRUCTURE: the figures go into BASELINE.md next to
the C port's; nothing on the GPU box runs this.

    python oracle/time_reference.py [n_moves_per_worker]"""
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_loader  # noqa: E402

AK3 = [1, 1, 1, -2, -2, -2, -2] + [0] * 29 + [1, 2, 1, -2, -1, -2] + [0] * 30


def _walk(args):
    """n lockstep-equivalent transitions: ACEnv.step along a random walk from AK(3), horizon 2000 with reset"""
    seed, n = args
    ref = ref_loader.load()
    rng = np.random.default_rng(seed)
    env = ref.ACEnv(ref.ACEnvConfig(initial_state=np.array(AK3, dtype=np.int8), horizon_length=2000))
    acts = rng.integers(0, 12, n)
    t0 = time.perf_counter()
    for a in acts:
        _, _, d, tr, _ = env.step(int(a))
        if d or tr:
            env.reset()
    return time.perf_counter() - t0


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    cores = os.cpu_count() or 1
    out = {"host_cores": cores, "numpy": np.__version__}
    dt = _walk((0, n))
    out["acenv_step_one_core_per_s"] = n / dt
    with mp.get_context("spawn").Pool(cores) as pool:
        pool.map(_walk, [(100 + k, 200) for k in range(cores)])  # import + warm-up in every worker
        t0 = time.perf_counter()
        pool.map(_walk, [(k, n) for k in range(cores)])
        wall = time.perf_counter() - t0
    out["acenv_step_pool_per_s"] = cores * n / wall
    out["pool_workers"] = cores
    ref = ref_loader.load()
    import contextlib
    import io

    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        ok, path = ref.bfs(np.array(AK3, dtype=np.int8), 100000)
    out["bfs_ak3_budget_1e5_s"] = time.perf_counter() - t0
    out["bfs_solved"] = bool(ok)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
