"""CPU, world_size 2, gloo: neighbourhood_sizes_sharded (instances partitioned over ranks, one all-gather of the sizes)
with the host-compiled graph engine returns the reference's ball sizes on every rank."""
import json
import os
import sys

import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, rows, out):
    for p in (os.path.join(ROOT, "ac-solver-caltech_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from acx import graph
    from host_harness import harness

    res = {}
    for classic in (True, False):
        res[classic] = graph.neighbourhood_sizes_sharded(rows, 5, classic, device="cpu", instances_per_batch=2, chunk=20000,
                                                         _backend=harness.HostGraphBackend(rank)).tolist()
    out.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_ball_sizes_world2(golden_dir):
    with open(os.path.join(golden_dir, "barcode_golden.json")) as f:
        g = json.load(f)["neighbourhoods"]
    pick = [0, 1, 2, 5, 9]  # an odd count: the ranks get 3 and 2 presentations
    rows = [g["rows"][i] for i in pick]
    from host_harness import harness

    harness.build()  # once, before the ranks race for it
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29950 + os.getpid() % 40
    procs = [ctx.Process(target=_worker, args=(r, 2, port, rows, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        rank, res = q.get(timeout=600)
        got[rank] = res
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank in (0, 1):
        assert got[rank][True] == [g["classic"][i] for i in pick]
        assert got[rank][False] == [g["prime"][i] for i in pick]
