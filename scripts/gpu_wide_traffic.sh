#!/bin/bash
# DRAM traffic of the wide ONF path's kernels over two planner steps of the gridmap-wide workload (ncu, one lane)
mkdir -p gpurun_out
B="python bench.py --workload gridmap-wide --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve --no-extra"
timeout 300 $B > gpurun_out/plain_wide.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:wide_ -s 40 -c 40 --csv --log-file gpurun_out/wide_traffic.csv $B > gpurun_out/wide_traffic.log 2>&1
echo rc=$?
