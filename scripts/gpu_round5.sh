#!/bin/bash
# end-of-round GPU visit: the three workload lines + the reference arm + the ONF sweep
set -x
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload forest --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_forest.json 2> gpurun_out/bench_forest.err; echo "forest rc=$?"
timeout 900 python bench.py --workload gridmap-wide --steps 50 --warmup 5 --cpu-seconds 6 --no-onf-microbench > gpurun_out/bench_wide.json 2> gpurun_out/bench_wide.err; echo "wide rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 python scripts/onf_sweep.py --json gpurun_out/onf_sweep.json > gpurun_out/onf_sweep.log 2>&1; echo "sweep rc=$?"
tail -c 600 gpurun_out/bench.err
