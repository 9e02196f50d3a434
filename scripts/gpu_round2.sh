#!/bin/bash
# GPU visit: full parity suite, the three planner workloads, the ONF sweep.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload forest --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_forest.json 2> gpurun_out/bench_forest.err; echo "forest rc=$?"
timeout 900 python bench.py --workload gridmap-wide --steps 50 --warmup 5 --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_wide.json 2> gpurun_out/bench_wide.err; echo "wide rc=$?"
tail -3 gpurun_out/bench*.err
timeout 900 python scripts/onf_sweep.py --json gpurun_out/onf_sweep.json > gpurun_out/onf_sweep.log 2>&1; echo "sweep rc=$?"
tail -5 gpurun_out/onf_sweep.log
