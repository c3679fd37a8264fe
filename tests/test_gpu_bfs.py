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
        cases = json.load(f)["bfs"]  # every width, odd max_relator_length included (AK(2) at 7)
    assert len(cases) > 100 and any((len(c["row"]) // 2) % 2 for c in cases)
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

    from acx.bfs_sharded import GpuShard, PlaneShard, ShardedBFS

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
                stats = {}
                ook, opath = oracle.bfs(row, budget, False, cyc, stats=stats)
                for kind in (GpuShard, PlaneShard):
                    shard = kind(len(row) // 2, cyc, budget + 64, chunk, str(cuda))
                    if kind is PlaneShard and not shard.accepts(row):
                        continue
                    eng = ShardedBFS(row, budget, shard, chunk=chunk)
                    ok, path = eng.run()
                    assert (ok, path) == (ook, opath), (int(k), budget, cyc, kind.__name__)
                    assert eng.n_nodes == stats["n_seen"] and eng.n_expanded == stats["n_expanded"], kind.__name__
    finally:
        if created:
            dist.destroy_process_group()


def test_plane_nodes_pack_expand_unpack(cuda, oracle, golden_dir):
    """plane nodes: pack / unpack round trip, and acx_pbfs_expand == the oracle's 12 ACMoves (children equal to
    their parent are dropped, everything else arrives with the right order number, lengths and letters)"""
    import ctypes

    import torch

    from acx import _lib
    from acx.batch import _ptr
    from acx.bfs_sharded import PlaneShard
    from helpers import random_rows

    rng = np.random.default_rng(77)
    for m, cyc in ((36, False), (36, True), (18, True), (7, False), (25, True)):
        rows = random_rows(rng, 3000, m, kinds=(0, 1, 2, 3, 3, 2))
        sh = PlaneShard(m, cyc, 16, 16, str(cuda))
        nodes, ok = sh.pack(rows)
        ok = ok.cpu().numpy().astype(bool)
        assert ok.sum() > 100
        back = sh.unpack(nodes).cpu().numpy()
        assert (back[ok] == rows[ok]).all()
        good = nodes[torch.from_numpy(ok).to(cuda)].contiguous()
        rows_ok = rows[ok]
        f = good.shape[0]
        loc = torch.arange(f, device=cuda)
        gids = torch.arange(f, device=cuda) + 1000
        cursor = torch.zeros((2,), dtype=torch.int64, device=cuda)
        out = torch.zeros((12 * f, 4), dtype=torch.int64, device=cuda)
        dst = (ctypes.c_void_p * 1)(out.data_ptr())
        _lib.check(sh.L.acx_pbfs_expand(_ptr(good), _ptr(loc), _ptr(gids), 1000, f, m, int(cyc), 1, 12 * f, _ptr(cursor), dst,
                                        0, sh.stream), "acx_pbfs_expand")
        n = int(cursor[0])
        assert int(cursor[1]) == 0
        rec = out[:n]
        order = (((rec[:, 2] >> 36) & 0xFFFFFFF) | (((rec[:, 3] >> 36) & 0xFFFFF) << 28)).cpu().numpy()
        status = ((rec[:, 3] >> 56) & 3).cpu().numpy()
        state = rec.clone()
        state[:, 2] &= (1 << 36) - 1
        state[:, 3] &= (1 << 36) - 1
        got_rows = sh.unpack(state).cpu().numpy()
        ch, ln, st = oracle.expand12_batch(rows_ok, cyclical=cyc)
        ch, st = ch.reshape(f * 12, 2 * m), st.reshape(f * 12)
        assert len(set(order.tolist())) == n
        seen = np.zeros(f * 12, dtype=bool)
        seen[order] = True
        assert (got_rows == ch[order]).all()
        assert (status == st[order]).all()
        same = (ch == np.repeat(rows_ok, 12, axis=0)).all(axis=1) & (st == 0)
        assert (seen == ~same).all()  # exactly the children that differ from their parent (or raise) are sent


def test_plane_bfs_device_driven_single_rank(cuda, oracle, golden_dir):
    """acx.pbfs.PlaneBFS (device-driven rounds, owner-sorted line-sized runs, fused winner resolution, grandparent
    filter, on-device budget / success decision) against the oracle: (solved, path, len(tree_nodes), expanded) for small
    and large budgets, several chunk sizes, both reduction modes, solved and budget-cut exits"""
    import torch.distributed as dist

    from acx.pbfs import PlaneBFS, make_plane_shard
    from helpers import AK2, AK3, pad_row, repad

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29700 + os.getpid() % 200))
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=cuda)
    try:
        rows = dataset_rows(golden_dir)
        rng = np.random.default_rng(33)
        cases = [(repad(AK2, 8), 4000, False, 64), (repad(AK2, 8), 4000, False, 1 << 16), (pad_row(*AK3, 36), 200000, False, 4096),
                 (pad_row(*AK3, 36), 50000, True, 100000), (pad_row(*AK3, 36), 1500000, False, 1 << 18),
                 # levels of more than 64 rounds (grouped block by block from the round log), a chunk that is no power of two
                 (pad_row(*AK3, 36), 60000, False, 64), (pad_row(*AK3, 36), 40000, True, 100)]
        for k in rng.choice(len(rows), size=10, replace=False):
            for budget, chunk in ((1, 4), (13, 4), (14, 4), (700, 64), (30000, 256), (30000, 4096)):
                cases.append((rows[int(k)], budget, bool(rng.integers(0, 2)), chunk))
        n_solved = 0
        for row, budget, cyc, chunk in cases:
            row = np.asarray(row, dtype=np.int8)
            stats = {}
            ook, opath = oracle.bfs(row, budget, False, cyc, stats=stats)
            shard = make_plane_shard(row, cyc, budget + 64, chunk, str(cuda))
            assert shard is not None
            eng = PlaneBFS(row, budget, shard, chunk=chunk)
            ok, path = eng.run()
            assert (ok, path) == (ook, opath), (row.tolist(), budget, cyc, chunk)
            assert eng.n_nodes == stats["n_seen"] and eng.n_expanded == stats["n_expanded"], (budget, cyc, chunk)
            n_solved += int(ok)
        assert n_solved >= 1
    finally:
        if created:
            dist.destroy_process_group()


def test_window_partition_blocks_and_holes(cuda):
    """acx_bfs_window_partition groups local nodes by round; ids outside [lo, lo + n_windows * chunk) -- another block of
    the level's windows, or -1 = a hole of the node store -- are left out (numpy restatement)"""
    import torch

    from acx import _lib
    from acx.batch import _ptr

    L = _lib.lib()
    rng = np.random.default_rng(5)
    stream = torch.cuda.current_stream(cuda).cuda_stream
    for n, chunk, lo, nwin in ((1, 4, 0, 1), (5000, 64, 100, 7), (300000, 1000, 12345, 40), (300000, 1 << 12, 4096, 64), (70000, 3, 10, 2000)):
        g = rng.integers(lo - 3 * chunk, lo + (nwin + 5) * chunk, size=n)
        g[rng.random(n) < 0.05] = -1
        first = 17
        gid = torch.from_numpy(g).to(cuda)
        counts = torch.zeros(nwin, dtype=torch.int64, device=cuda)
        _lib.check(L.acx_bfs_window_partition(_ptr(gid), n, lo, chunk, nwin, _ptr(counts), 0, 0, first, 0, 0, 0, stream), "count")
        keep = (g >= lo) & (g < lo + nwin * chunk)
        w = (g[keep] - lo) // chunk
        want = np.bincount(w, minlength=nwin)
        assert counts.cpu().numpy().tolist() == want.tolist(), (n, chunk)
        ends = torch.cumsum(counts, 0)
        offsets = (ends - counts).contiguous()
        cursor = torch.zeros(nwin, dtype=torch.int64, device=cuda)
        loc = torch.full((n,), -7, dtype=torch.int64, device=cuda)
        out = torch.full((n,), -7, dtype=torch.int64, device=cuda)
        _lib.check(L.acx_bfs_window_partition(_ptr(gid), n, lo, chunk, nwin, _ptr(counts), _ptr(offsets), _ptr(cursor), first,
                                              _ptr(loc), _ptr(out), 1, stream), "scatter")
        loc, out, e = loc.cpu().numpy(), out.cpu().numpy(), ends.cpu().numpy()
        total = int(keep.sum())
        assert (loc[total:] == -7).all() and (out[total:] == -7).all()
        assert (g[loc[:total] - first] == out[:total]).all()                  # every entry names a node and carries its id
        assert sorted(loc[:total].tolist()) == (np.flatnonzero(keep) + first).tolist()  # each kept node exactly once
        b = 0
        for k in range(nwin):                                                   # grouped by window, any order inside
            seg = out[b:e[k]]
            assert ((seg - lo) // chunk == k).all()
            b = e[k]


def test_plane_bfs_store_invariants(cuda, oracle):
    """after a search that ran into its budget: every place of the node store below the last round is a node with a
    global id, or a hole (gid -1, zero node) that no table slot names; ids 0 .. n-1 appear exactly once"""
    import torch
    import torch.distributed as dist

    from acx.pbfs import PlaneBFS, make_plane_shard
    from helpers import AK3, pad_row

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29700 + os.getpid() % 200))
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=cuda)
    try:
        row, budget = pad_row(*AK3, 36), 300000
        shard = make_plane_shard(row, False, budget + 64, 1 << 12, str(cuda))
        eng = PlaneBFS(row, budget, shard, chunk=1 << 12)
        ok, _ = eng.run()
        assert not ok
        n_local = int(shard.n_local_dev.item())
        gid = shard.gid[:n_local].cpu().numpy()
        nodes = shard.nodes[:n_local].cpu().numpy()
        committed = (gid >= 0) & (gid < (1 << 62))
        holes = gid == -1
        tail = ~committed & ~holes                      # the round the budget cut (and the one behind it): orders, not ids
        assert (nodes[holes] == 0).all() and (nodes[committed, 0] != 0).all()
        ids = np.sort(gid[committed])
        assert (ids == np.arange(len(ids))).all()        # the discovery ranks of the reference, each exactly once
        assert len(ids) <= eng.n_nodes <= len(ids) + int(tail.sum())
        assert holes.sum() <= 0.1 * n_local + 4096
    finally:
        if created:
            dist.destroy_process_group()
