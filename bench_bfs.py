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


if __name__ == "__main__":
    import json

    b = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10 ** 7
    print(json.dumps(run(torch.device("cuda:0"), b)))
