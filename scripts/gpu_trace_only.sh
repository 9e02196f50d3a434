#!/bin/bash
# quick parity run of the tensor path, phase trace of the streamed ONF kernel, short bench lines
timeout 600 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-onf or lockstep or warmup or free_run or api}" 2>&1 | tail -8
mkdir -p gpurun_out
for dbg in ${DBGS:-0}; do
NFOPP_TCS_DBG=$dbg NFOPP_LANES=1 timeout 300 python scripts/tcs_trace.py > gpurun_out/tcs_trace_dbg$dbg.txt 2>&1; echo trace dbg=$dbg rc=$?
done
LANES="${LANES:-1 4}" DBGS=0 scripts/gpu_dbg_sweep.sh
