#!/usr/bin/env python
"""Summary of a PlaneBFS pipeline trace (ACX_BFS_TRACE=<prefix>; acx/pbfs.py): where the wall time of a search goes.

    python tools/bfs_trace_summary.py gpurun_out/bfs_trace_rank0.json [first round to list] [rounds to list]"""
import collections
import json
import sys

rows = json.load(open(sys.argv[1]))
ev = collections.defaultdict(dict)
for name, k, t in rows:
    ev[k][name] = t
rounds = sorted(k for k in ev if "probe_begin" in ev[k] and "probe_end" in ev[k])
total = max(t for _, _, t in rows)
dur = lambda k, a, b: ev[k][b] - ev[k][a] if a in ev[k] and b in ev[k] else 0.0
probe = sum(dur(k, "probe_begin", "probe_end") for k in rounds)
expand = sum(dur(k, "expand_begin", "expand_end") for k in ev)
gather = sum(dur(k, "allgather_begin", "allgather_end") for k in ev)
red = sum(dur(k, "reduce_begin", "reduce_end") for k in ev)
com = sum(dur(k, "reduce_end", "commit_end") for k in ev)
gaps = sum(max(0.0, ev[b]["probe_begin"] - ev[a]["probe_end"]) for a, b in zip(rounds, rounds[1:]))
print("search %.1f us, %d rounds; in-kernel spans (they overlap across streams): probe %.1f, expand %.1f, count all-gather %.1f, "
      "mask/min all-reduce %.1f, scan+decide+commit %.1f; main stream idle between probes %.1f" % (total, len(rounds), probe, expand, gather, red, com, gaps))
first = int(sys.argv[2]) if len(sys.argv) > 2 else (rounds[len(rounds) // 2] if rounds else 0)
n = int(sys.argv[3]) if len(sys.argv) > 3 else 6
print("round: expand [begin..end]  all-gather [begin..end]  probe [begin..end]  reduce [begin..end]  commit end   (us since the start)")
for k in rounds:
    if first <= k < first + n:
        g = lambda name: ev[k].get(name, float("nan"))
        print("%5d: %9.1f %9.1f | %9.1f %9.1f | %9.1f %9.1f | %9.1f %9.1f | %9.1f" % (k, g("expand_begin"), g("expand_end"), g("allgather_begin"),
              g("allgather_end"), g("probe_begin"), g("probe_end"), g("reduce_begin"), g("reduce_end"), g("commit_end")))
