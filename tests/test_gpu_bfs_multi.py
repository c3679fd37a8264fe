"""GPU, world_size 2 (NCCL): the hash-partitioned bfs with the product shard backend (libacx kernels, fused
expand+bucket, all-to-all over NVLink) returns exactly the oracle's (solved, path, len(tree_nodes), expanded) on
every rank.  Skipped on a one-GPU box (the one-rank variant is in test_gpu_bfs.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _worker(rank, world, port, cases, out):
    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from acx.bfs_sharded import GpuShard, PlaneShard, ShardedBFS

    res = []
    for row, budget, cyc, chunk in cases:
        row = np.array(row, dtype=np.int8)
        for kind, transport in ((GpuShard, "auto"), (GpuShard, "nccl"), (PlaneShard, "auto"), (PlaneShard, "nccl")):
            shard = kind(len(row) // 2, cyc, budget + 64, chunk, str(dev), transport=transport)  # auto = peer stores over NVLink
            if kind is PlaneShard and not shard.accepts(row):
                res.append(None)
                continue
            eng = ShardedBFS(row, budget, shard, chunk=chunk, pipeline=(transport == "auto"))  # both schedules
            ok, path = eng.run()
            shard.close()
            res.append((ok, path, eng.n_nodes, eng.n_expanded, shard.transport))
        # the device-driven engine (acx.pbfs) on the same case
        from acx.pbfs import PlaneBFS, make_plane_shard

        shard = make_plane_shard(row, cyc, budget + 64, chunk, str(dev))
        if shard is None:
            res.append(None)
        else:
            eng = PlaneBFS(row, budget, shard, chunk=chunk)
            ok, path = eng.run()
            shard.close()
            res.append((ok, path, eng.n_nodes, eng.n_expanded, shard.transport))
    out.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_bfs_two_gpus_matches_oracle(cuda, oracle, golden_dir):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from helpers import AK2, AK3, dataset_rows, pad_row, repad

    rows = [r for r in dataset_rows(golden_dir)]
    rng = np.random.default_rng(9)
    cases = [(repad(AK2, 8).tolist(), 4000, False, 64), (pad_row(*AK3, 36).tolist(), 200000, False, 4096),
             (pad_row(*AK3, 36).tolist(), 50000, True, 100000)]
    for k in rng.choice(len(rows), size=6, replace=False):
        for budget, chunk in ((1, 8), (13, 8), (900, 32), (30000, 2048)):
            cases.append((rows[int(k)].tolist(), budget, bool(rng.integers(0, 2)), chunk))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29800 + os.getpid() % 150
    procs = [ctx.Process(target=_worker, args=(r, 2, port, cases, q)) for r in range(2)]
    for p in procs:
        p.start()
    import queue as _queue
    import time as _time

    got, t_end = {}, _time.time() + 600
    while len(got) < 2:
        try:
            rank, res = q.get(timeout=2)
            got[rank] = res
        except _queue.Empty:
            assert all(p.exitcode in (None, 0) for p in procs), "a rank died: %s" % [p.exitcode for p in procs]
            assert _time.time() < t_end, "timed out"
    for p in procs:
        p.join(timeout=60)
    for idx, (row, budget, cyc, chunk) in enumerate(cases):
        stats = {}
        ok, path = oracle.bfs(np.array(row, dtype=np.int8), budget, False, cyc, stats=stats)
        for rank in (0, 1):
            for t in range(5):
                if got[rank][5 * idx + t] is None:
                    continue
                rok, rpath, n_seen, n_exp, transport = got[rank][5 * idx + t]
                assert (rok, rpath) == (ok, path), (idx, rank, t, transport)
                assert n_seen == stats["n_seen"] and n_exp == stats["n_expanded"], (idx, rank, t, transport)
        assert got[0][5 * idx][4] == ("p2p" if len(row) == 72 else "nccl")
        assert got[0][5 * idx + 2] is None or got[0][5 * idx + 2][4] == "p2p"
