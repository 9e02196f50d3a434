#!/bin/bash
# One GPU visit: gpu tests (optional filter), a short bench line, and (optionally) ncu of one kernel.  Parametrised successor of
# the round-1 gpu_round*.sh scripts.   usage: scripts/gpu_job.sh [tests|bench|ncu-launches|ncu-full <regex> <skip>|all] ...
set -x
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve --no-extra"
for job in "$@"; do
case "$job" in
  tests) timeout 1500 python -m pytest tests -m gpu -q -x ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log;;
  tests-all) timeout 1500 python -m pytest tests -m gpu -q ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_gpu.log;;
  bench) timeout 900 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err;;
  bench-quick) timeout 600 python bench.py --steps 100 --warmup 10 --no-extra --no-cpu-baseline --no-onf-microbench > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_quick.err;;
  ncu-launches) timeout 300 $BENCH > gpurun_out/plain1.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?";;
  ncu-full) timeout 300 $BENCH > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:${NCU_K:-onf_tcs_kernel} -s ${NCU_S:-6} -c ${NCU_C:-3} -f -o gpurun_out/${NCU_O:-onf_tcs_full} $BENCH > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log;;
esac
done
ls -la gpurun_out | head -30
