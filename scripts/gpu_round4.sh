#!/bin/bash
# GPU visit for the committed round numbers: parity suite, three workload lines, config-5 sweep, ncu launch lists and
# full captures of the dominant kernels (corridor: onf_tc train launch; wide: chain + wgrad2).
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload forest --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_forest.json 2> gpurun_out/bench_forest.err; echo "forest rc=$?"
timeout 900 python bench.py --workload gridmap-wide --steps 50 --warmup 5 --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_wide.json 2> gpurun_out/bench_wide.err; echo "wide rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 python scripts/onf_sweep.py --json gpurun_out/onf_sweep.json > gpurun_out/onf_sweep.log 2>&1; echo "sweep rc=$?"
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
ls -la gpurun_out | head -40
