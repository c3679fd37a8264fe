#!/bin/bash
# plane bfs on N GPUs (torchrun): one run per argument "VAR=value ..."; usage: bash tools/bfs_sweep_multi.sh N "ACX_BFS_CHUNK=2097152" ...
N=$1; shift
export ACX_BFS_LARGE=0
for cfg in "$@"; do
  echo "# N=$N $cfg"
  env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench_bfs.py sharded 1e9 2>gpurun_out/bfs_multi.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(json.dumps({k:d[k] for k in ('nodes_expanded_per_s','wall_s','parity','phase_ms','rounds','store_holes','nvlink_egress_gbs_per_rank')}))
" || tail -5 gpurun_out/bfs_multi.err
done
