"""BFS nodes expanded / s (BASELINE metric, configs[3] on one GPU): level-synchronous bfs from AK(3), m=36."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))


def ak3(m=36):
    row = np.zeros(2 * m, dtype=np.int8)
    row[:7] = [1, 1, 1, -2, -2, -2, -2]
    row[m:m + 6] = [1, 2, 1, -2, -1, -2]
    return row


def run(dev, budget=10 ** 7, chunk=1 << 20):
    from acx.bfs import LevelBFS

    LevelBFS(ak3(), 10 ** 5, False, str(dev), chunk).run()  # warm-up
    torch.cuda.synchronize()
    eng = LevelBFS(ak3(), budget, False, str(dev), chunk)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ok, path = eng.run()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"config": "configs[3] level-synchronous bfs from AK(3), max_relator_length=36, budget %d (1 GPU)" % budget,
            "solved": ok, "n_seen": eng.n_nodes, "nodes_expanded": eng.n_expanded, "wall_s": dt,
            "nodes_expanded_per_s": eng.n_expanded / dt, "discovered_per_s": eng.n_nodes / dt, "levels": eng.levels,
            "reference": "660 expanded nodes/s, 5.6k discovered nodes/s at budget 1e5 (SURVEY.md section 6)"}


def run_sharded(budget=10 ** 8, chunk_per_rank=1 << 20):
    """torchrun entry: hash-partitioned bfs over WORLD_SIZE GPUs, strong scaling at a fixed budget.
    A round expands `chunk_per_rank` parents per rank (the result does not depend on the chunk)."""
    import torch.distributed as dist

    from acx.bfs_sharded import GpuShard, ShardedBFS, make_shard

    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ["NCCL_DEBUG"] = "WARN"
    dist.init_process_group("nccl", device_id=dev)
    out = None
    chunk = chunk_per_rank * world
    if world > 1:  # NCCL connects lazily per algorithm / message size: touch the sizes the search uses
        for _ in range(3):
            dist.all_reduce(torch.zeros(chunk, dtype=torch.int32, device=dev))
            dist.all_reduce(torch.zeros(2, dtype=torch.int64, device=dev), op=dist.ReduceOp.MIN)
            dist.all_gather_into_tensor(torch.zeros(world * (world + 1), dtype=torch.int64, device=dev),
                                        torch.zeros(world + 1, dtype=torch.int64, device=dev))
    for b in (10 ** 6, budget):  # first pass = warm-up (NCCL channels, kernels)
        cap_local = int(b / world * (1.0 if world == 1 else 1.06)) + 4096
        max_parents = chunk if world == 1 else min(chunk, int(chunk / world * 1.15) + 8192)
        transport = os.environ.get("ACX_BFS_TRANSPORT", "auto")
        if os.environ.get("ACX_BFS_NODES", "planes") == "bytes":
            shard = GpuShard(36, False, cap_local, max_parents, str(dev), transport=transport)
        else:
            shard = make_shard(ak3(), False, cap_local, max_parents, str(dev), transport=transport)
        eng = ShardedBFS(ak3(), b, shard, chunk=chunk, pipeline=os.environ.get("ACX_BFS_PIPELINE", "0") != "0",
                         stage_ctas=int(os.environ.get("ACX_BFS_STAGE_CTAS", "2")))
        if os.environ.get("ACX_BFS_PROF"):
            eng.timers = {}
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ok, path = eng.run()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        sent = torch.tensor([eng.comm_bytes], device=dev, dtype=torch.int64)
        dist.all_reduce(sent)
        out = {"config": "configs[3] hash-partitioned level-synchronous bfs from AK(3), m=36, budget %d" % b,
               "n_gpus": world, "solved": ok, "n_seen": eng.n_nodes, "nodes_expanded": eng.n_expanded, "wall_s": float(dt[0]),
               "nodes_expanded_per_s": eng.n_expanded / float(dt[0]), "all_to_all_bytes_total": int(sent[0]),
               "scaling": "strong", "chunk_parents_per_round": chunk, "transport": shard.transport, "nodes": type(shard).__name__, "levels": eng.levels}
        if eng.timers is not None:
            out["phase_s_rank0"] = {k: round(v, 5) for k, v in eng.timers.items()}
        shard.close()
        del eng, shard
        torch.cuda.empty_cache()
    if rank == 0:
        import json
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "sharded":
    run_sharded(int(float(sys.argv[2])) if len(sys.argv) > 2 else 10 ** 8,
                int(float(sys.argv[3])) if len(sys.argv) > 3 else 1 << 20)
elif __name__ == "__main__":
    import json

    b = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10 ** 7
    print(json.dumps(run(torch.device("cuda:0"), b)))
