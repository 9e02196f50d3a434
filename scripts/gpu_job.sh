#!/bin/bash
# One GPU visit, any number of jobs in order.   usage: [VAR=...] scripts/gpu_job.sh <job> [<job> ...]
#   tests | tests-all      GPU test suite (-x / all), optional filter PYTEST_K
#   tests-debug            the same against libnfopp_debug.so (device-side bounds asserts; make -C csrc debug first)
#   smoke                  __graft_entry__.smoke()
#   bench | bench-quick    full / reduced bench line -> gpurun_out/bench*.json
#   bench-ref              the CPU arm (--impl reference)
#   sweep                  corridor bench lines for LANES="1 4" x DBGS="0" (NFOPP_TCS_DBG switches), WORKLOAD / PROBLEMS optional
#   kernels                CUDA-event times of the step's kernels (one lane), twice
#   trace                  phase trace of the streamed ONF kernel (make -C csrc trace first) -> gpurun_out/tcs_trace.txt
#   single                 single-problem latency through the planner API (configs[0])
#   soak                   scripts/soak.py (5000 steps per workload)
#   ncu-launches           launch list (gpu__time_duration) of the short bench -> gpurun_out/launches.csv
#   ncu-full               ncu --set full of NCU_K (regex), skip NCU_S, count NCU_C -> gpurun_out/$NCU_O.ncu-rep
#   ncu-wide-traffic       DRAM bytes of the wide path's kernels (gridmap-wide) -> gpurun_out/wide_traffic.csv
# ncu jobs run the plain command first and profile only if it exits 0.
set -x
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve --no-extra"
QUICK="--no-extra --no-cpu-baseline --no-onf-microbench --no-solve"
line() {   # one summary line of a bench JSON file
  python - "$1" "$2" <<'PY'
import json, sys
d = json.load(open(sys.argv[1])); k = d["kernel_ms"]
print("%s value %.0f ms/step %.4f e2e %.0f | cand %.3f train %.3f path %.3f | sample_pre %.4f resample %.4f traj_post %.4f traj_pre %.4f tick %.4f"
      % (sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], k["onf_forward_candidates"], k["onf_train_fwd_bwd"],
         k["onf_path_fwd_dgrad"], k["sample_pre"], k["resample_label"], k["traj_post"], k["traj_pre"], k["tick"]))
PY
}
for job in "$@"; do
case "$job" in
  tests) timeout 1500 python -m pytest tests -m gpu -q -x ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log;;
  tests-all) timeout 1500 python -m pytest tests -m gpu -q ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_gpu.log;;
  tests-debug) NFOPP_LIB=libnfopp_debug.so timeout 1500 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-not labels_large}" > gpurun_out/pytest_debug.log 2>&1; echo "pytest(debug) rc=$?"; tail -5 gpurun_out/pytest_debug.log;;
  smoke) python __graft_entry__.py --smoke 2>&1 | tail -2;;
  bench) timeout 900 python bench.py --steps 100 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err;;
  bench-quick) timeout 600 python bench.py --steps 100 --warmup 10 --no-extra --no-cpu-baseline --no-onf-microbench > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_quick.err;;
  bench-ref) timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench-ref rc=$?";;
  sweep) for dbg in ${DBGS:-0}; do for lanes in ${LANES:-1 4}; do
           NFOPP_TCS_DBG=$dbg NFOPP_LANES=$lanes timeout 600 python bench.py --workload ${WORKLOAD:-corridor} ${PROBLEMS:+--problems $PROBLEMS} --steps ${STEPS:-100} --warmup 10 $QUICK > /tmp/b.json && line /tmp/b.json "dbg$dbg lanes$lanes"
         done; done;;
  kernels) for i in 1 2; do NFOPP_LANES=1 timeout 300 python bench.py --steps 100 --warmup 10 $QUICK > /tmp/b.json && line /tmp/b.json "lanes1"; done;;
  trace) NFOPP_LANES=1 timeout 300 python scripts/tcs_trace.py > gpurun_out/tcs_trace.txt 2>&1; echo "trace rc=$?";;
  single) python -c "import bench, json; print(json.dumps(bench.single_problem_bench()))";;
  soak) timeout 1200 python scripts/soak.py 2>&1 | tail -5;;
  ncu-launches) timeout 300 $BENCH > gpurun_out/plain1.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?";;
  ncu-full) timeout 300 $BENCH > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:${NCU_K:-onf_tcs_kernel} -s ${NCU_S:-6} -c ${NCU_C:-3} -f -o gpurun_out/${NCU_O:-onf_tcs_full} $BENCH > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log;;
  ncu-wide-traffic) W="python bench.py --workload gridmap-wide --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve --no-extra"
         timeout 300 $W > gpurun_out/plain_wide.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:wide_ -s 40 -c 40 --csv --log-file gpurun_out/wide_traffic.csv $W > gpurun_out/wide_traffic.log 2>&1; echo "ncu wide traffic rc=$?";;
  *) echo "unknown job $job";;
esac
done
ls -la gpurun_out | head -30
