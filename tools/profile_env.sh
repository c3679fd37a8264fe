#!/bin/bash
# ncu evidence for the env-step kernel (run under gpurun, one GPU): launch list, full-set capture of two timed launches,
# and the DRAM traffic of 8 consecutive launches without cache control (steady state: the dirty lines of launch k-1 are
# written back while launch k runs).  Outputs land in gpurun_out/.
set -x
CMD="python bench.py --steps 8 --warmup 3 --no-graph --no-cpu-baseline --no-bfs --no-other"
export ACX_BENCH_CHAINED=0
$CMD > gpurun_out/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 60 --csv --log-file gpurun_out/r02_launches_bench.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_env_tiles -s 40 -c 2 -o gpurun_out/r02_k_env_tiles $CMD > gpurun_out/ncu2.log 2>&1
ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:k_env_tiles -s 40 -c 8 --csv --log-file gpurun_out/r02_env_traffic.csv $CMD > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
