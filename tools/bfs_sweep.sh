#!/bin/bash
# tuning sweep of the plane-bfs round kernels on one GPU (probe occupancy, stream overlap)
# usage (GPU box): bash tools/bfs_sweep.sh [VAR=value ...]... > gpurun_out/bfs_sweep.jsonl    (one run per argument)
export ACX_BFS_LARGE=0
run() {
  echo "# $*"
  env $* python bench_bfs.py 1e9 2>gpurun_out/bfs_sweep.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({k:d[k] for k in ('nodes_expanded_per_s','wall_s','parity','phase_ms','rounds','store_holes')}))
" || tail -5 gpurun_out/bfs_sweep.err
}
if [ $# -eq 0 ]; then set -- "ACX_PROBE_OCC=6" "ACX_PROBE_OCC=4" "ACX_PROBE_OCC=8" "ACX_PROBE_OCC=6 ACX_BFS_OVERLAP=0"; fi
for cfg in "$@"; do run $cfg; done
