#!/usr/bin/env python
"""bench_graph.py -- timing / profiling driver of the graph engine (ac_bfs.cpp).

    python bench_graph.py [n=14] [classic|prime] [chunk=1048576] [mode]

Prints one JSON line: nodes, edges, per-level frontier sizes and device times (CUDA events around every level), the time of
the whole BFS and of the BFS plus the edge list on the host.  mode -2 runs a second pass with the CUDA profiler bracket
around the largest level only:  ncu --set full --clock-control none --profile-from-start off python bench_graph.py 14 classic 1048576 -2
(the captures summarised in profiles/r01_graph_n14_largest_level_ncu*.txt)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ac-solver-caltech_b200"))
import numpy as np
import torch
from acx import graph

n = int(sys.argv[1]) if len(sys.argv) > 1 else 14
moves = sys.argv[2] if len(sys.argv) > 2 else "classic"
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
profile_level = int(sys.argv[4]) if len(sys.argv) > 4 else -1


def run(profile_level):
    eng = graph.GraphEngine(2, moves == "classic", max_total=n, want_edges=True, chunk=chunk, device="cuda:0")
    eng.seed(np.array([[1, 2]], dtype=np.int8))
    lo, hi, levels, times = 0, 1, [], []
    torch.cuda.synchronize()
    t_all = time.perf_counter()
    while lo < hi:
        levels.append(hi - lo)
        if len(levels) - 1 == profile_level:
            torch.cuda.profiler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.expand_level(lo, hi)
        e1.record()
        torch.cuda.synchronize()
        if len(levels) - 1 == profile_level:
            torch.cuda.profiler.stop()
        times.append(e0.elapsed_time(e1))
        lo, hi = hi, eng.n_nodes
    t_bfs = time.perf_counter() - t_all
    edges = torch.cat(eng.edge_parts).cpu()
    filt = torch.cat(eng.filt_parts).cpu()
    sizes = eng.node_size[: eng.n_nodes].cpu()
    t_total = time.perf_counter() - t_all
    return eng, levels, times, t_bfs, t_total, len(edges)


run(-1) if n <= 12 else graph.ac_graph_bfs(8, moves)  # warm-up
eng, levels, times, t_bfs, t_total, ne = run(-1)
big = int(np.argmax(levels))
nm = eng.nm
print(json.dumps({"n": n, "moves": moves, "chunk": chunk, "nodes": eng.n_nodes, "edges": ne, "levels": len(levels),
                  "bfs_s": t_bfs, "with_d2h_s": t_total, "device_ms_sum": sum(times),
                  "largest_level": {"index": big, "parents": levels[big], "candidates": levels[big] * nm, "ms": times[big],
                                    "candidates_per_s": levels[big] * nm / (times[big] * 1e-3)},
                  "level_sizes": levels, "level_ms": [round(t, 3) for t in times]}))
if profile_level == -2:  # second pass with the profiler bracket around the largest level
    del eng
    torch.cuda.empty_cache()
    run(big)
