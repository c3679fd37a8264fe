"""GPU: the level-synchronous bfs (visited table with first-discovery semantics, scan-assigned ids, budget
rule on the scanned flags) returns exactly what the FIFO reference returns."""
import json
import os

import numpy as np
import pytest

from helpers import dataset_rows

pytestmark = pytest.mark.gpu


def test_bfs_level_golden(cuda, golden_dir):
    from acx.bfs import bfs_level

    with open(os.path.join(golden_dir, "search_golden.json")) as f:
        cases = [c for c in json.load(f)["bfs"] if (len(c["row"]) // 2) % 2 == 0]
    assert len(cases) > 100
    for k, c in enumerate(cases):
        stats = {}
        ok, path = bfs_level(np.array(c["row"], dtype=np.int8), c["budget"], c["cyclical"], str(cuda),
                             chunk=(64, 1000, 1 << 20)[k % 3], stats=stats)
        assert ok == c["solved"], c
        assert path == (None if c["path"] is None else [tuple(p) for p in c["path"]]), c
        if c["n_seen_at_exit"] is not None:
            assert stats["n_seen"] == c["n_seen_at_exit"], c


@pytest.mark.parametrize("chunk", [7, 256, 1 << 20])
def test_bfs_level_vs_oracle(cuda, oracle, golden_dir, chunk):
    from acx.bfs import bfs_level

    rows = dataset_rows(golden_dir)
    rng = np.random.default_rng(chunk)
    for k in rng.choice(len(rows), size=14, replace=False):
        for budget in (1, 2, 13, 14, 300, 5000, 60000):
            cyc = bool(rng.integers(0, 2))
            stats, ostats = {}, {}
            ok, path = bfs_level(rows[int(k)], budget, cyc, str(cuda), chunk=chunk, stats=stats)
            ook, opath = oracle.bfs(rows[int(k)], budget, False, cyc, stats=ostats)
            assert (ok, path) == (ook, opath), (int(k), budget, cyc)
            assert stats["n_seen"] == ostats["n_seen"] and stats["n_expanded"] == ostats["n_expanded"], (int(k), budget, cyc)


def test_bfs_level_ak3_levels(cuda):
    """AK(3) at m=36: new nodes per depth as measured on the reference, budget 1e5 -> unsolved"""
    from acx.bfs import bfs_level
    from helpers import AK3, pad_row

    stats = {}
    ok, path = bfs_level(pad_row(AK3[0], AK3[1], 36), 100000, False, str(cuda), stats=stats)
    assert not ok and path is None
    assert stats["levels"][:6] == [1, 12, 116, 1045, 9142, 76612]


def test_sharded_bfs_single_rank_gpu(cuda, oracle, golden_dir):
    """the product shard backend (GpuShard: libacx kernels) under the sharded driver, world_size 1 (NCCL)"""
    import torch
    import torch.distributed as dist

    from acx.bfs_sharded import GpuShard, ShardedBFS

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29700 + os.getpid() % 200))
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=cuda)
    try:
        rows = dataset_rows(golden_dir)
        rng = np.random.default_rng(21)
        for k in rng.choice(len(rows), size=8, replace=False):
            for budget, chunk in ((1, 4), (14, 4), (700, 64), (30000, 4096)):
                cyc = bool(rng.integers(0, 2))
                row = rows[int(k)]
                shard = GpuShard(len(row) // 2, cyc, budget + 64, chunk, str(cuda))
                eng = ShardedBFS(row, budget, shard, chunk=chunk)
                ok, path = eng.run()
                stats = {}
                ook, opath = oracle.bfs(row, budget, False, cyc, stats=stats)
                assert (ok, path) == (ook, opath), (int(k), budget, cyc)
                assert eng.n_nodes == stats["n_seen"] and eng.n_expanded == stats["n_expanded"]
    finally:
        if created:
            dist.destroy_process_group()
