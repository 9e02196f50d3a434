#!/bin/bash
# GPU experiment round for the persistent streamed ONF kernel: parity tests of the tensor path, then the corridor bench under
# several grid caps / lane counts, then the phase trace.   usage: scripts/gpu_exp.sh
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-onf or lockstep or warmup or free_run or independent or lanes or full_size_corridor}" > gpurun_out/pytest_exp.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/pytest_exp.log
Q="--steps 100 --warmup 10 --no-extra --no-cpu-baseline --no-onf-microbench --no-solve"
for cfg in ${EXP_CFGS:-"2:0" "1:0" "2:74" "1:148" "2:148" "1:1"}; do
  lanes=${cfg%%:*}; grid=${cfg##*:}
  if [ "$grid" = "0" ]; then unset NFOPP_TCS_GRID; else export NFOPP_TCS_GRID=$grid; fi
  [ "$grid" = "1" ] && export NFOPP_TCS_GRID=0
  NFOPP_LANES=$lanes timeout 300 python bench.py $Q > gpurun_out/bench_l${lanes}_g${grid}.json 2> gpurun_out/bench_l${lanes}_g${grid}.err
  echo "lanes=$lanes grid=$grid rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_l${lanes}_g${grid}.json"))
    k = d["kernel_ms"]
    print("  value %.0f ms/step %.4f e2e %.0f | cand %.3f train %.3f path %.3f label %.3f traj_post %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], k["onf_forward_candidates"], k["onf_train_fwd_bwd"], k["onf_path_fwd_dgrad"], k["resample_label"], k["traj_post"]))
except Exception as e:
    print("  failed:", e)
PY
done
unset NFOPP_TCS_GRID
if [ -n "$EXP_TRACE" ]; then
  # (libnfopp_trace.so is built in the container: make -C pytorch-motion-planner_b200/csrc trace)
  NFOPP_LANES=1 timeout 300 python scripts/tcs_trace.py > gpurun_out/tcs_trace.txt 2>&1; echo "trace rc=$?"; cat gpurun_out/tcs_trace.txt | cut -c1-600
fi
