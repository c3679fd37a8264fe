#!/bin/bash
# ncu evidence for the plane-bfs kernels (one GPU): the four kernels of the last full-size rounds of a budget-6e7 search
set -x
CMD="python tools/bfs_profile_run.py 6e7"
$CMD > gpurun_out/plain_bfs.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_pbfs_ -s 60 -c 8 -o gpurun_out/r02_pbfs6 $CMD > gpurun_out/ncu_bfs.log 2>&1
tail -2 gpurun_out/ncu_bfs.log
