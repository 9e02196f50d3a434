#!/bin/bash
# ncu evidence for the end-of-round code: launch lists + full captures of the dominant kernels (each only after the same
# command has exited 0 without ncu)
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve > gpurun_out/plain1.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:onf_tc -s 6 -c 3 -f -o gpurun_out/onf_tc_full \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 300 python scripts/onf_sweep.py --quick --reps 1 > gpurun_out/plain2.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_wide.csv \
  python scripts/onf_sweep.py --quick --reps 1 > gpurun_out/ncu_wide.log 2>&1; echo "ncu wide launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wide_chain|wide_wgrad2" -s 0 -c 2 -f -o gpurun_out/wide_full \
  python scripts/onf_sweep.py --quick --reps 1 > gpurun_out/ncu_wide_full.log 2>&1; echo "ncu wide full rc=$?"
ls -la gpurun_out | head -30
