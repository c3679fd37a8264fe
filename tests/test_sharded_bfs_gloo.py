"""CPU, world_size 2, gloo: the host side of the hash-partitioned bfs (acx.bfs_sharded.ShardedBFS) with the
oracle-backed stand-in shard gives exactly the oracle's (solved, path, len(tree_nodes)) on every rank."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, cases, out):
    for p in (os.path.join(ROOT, "ac-solver-caltech_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from acx.bfs_sharded import ShardedBFS
    from cpu_shard import OracleShard

    res = []
    for row, budget, cyc, chunk in cases:
        row = np.array(row, dtype=np.int8)
        eng = ShardedBFS(row, budget, OracleShard(len(row) // 2, cyc, budget + 64), chunk=chunk)
        ok, path = eng.run()
        res.append((ok, path, eng.n_nodes, eng.n_expanded))
    out.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_bfs_world2_matches_oracle(oracle, golden_dir):
    from helpers import AK2, dataset_rows

    rows = dataset_rows(golden_dir)
    rng = np.random.default_rng(4)
    cases = [(AK2.tolist(), 4000, False, 50)]
    for k in rng.choice(len(rows), size=6, replace=False):
        for budget, chunk in ((1, 8), (13, 8), (150, 16), (900, 1000)):
            cases.append((rows[int(k)].tolist(), budget, bool(rng.integers(0, 2)), chunk))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, cases, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    import queue as _queue
    import time as _time
    t_end = _time.time() + 600
    while len(got) < 2:
        try:
            rank, res = q.get(timeout=2)
            got[rank] = res
        except _queue.Empty:
            assert all(p.exitcode in (None, 0) for p in procs), "a rank died: %s" % [p.exitcode for p in procs]
            assert _time.time() < t_end, "timed out"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for idx, (row, budget, cyc, chunk) in enumerate(cases):
        stats = {}
        ok, path = oracle.bfs(np.array(row, dtype=np.int8), budget, False, cyc, stats=stats)
        for rank in (0, 1):
            rok, rpath, n_seen, n_exp = got[rank][idx]
            assert (rok, rpath) == (ok, path), (idx, rank)
            assert n_seen == stats["n_seen"] and n_exp == stats["n_expanded"], (idx, rank)


def test_popcount_and_owner():
    sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
    from acx.bfs_sharded import owner_of, popcount12

    x = torch.tensor([0, 1, 0xFFF, 0x555, 0x800], dtype=torch.int64)
    assert popcount12(x).tolist() == [0, 1, 12, 6, 1]
    h = torch.tensor([-1, 0, 1 << 40, -(1 << 62)], dtype=torch.int64)
    o = owner_of(h, 8)
    assert ((o >= 0) & (o < 8)).all()
