#!/bin/bash
# quick parity run of the tensor path, phase trace of the streamed ONF kernel (train launch), short bench lines
timeout 600 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-onf or lockstep or warmup or free_run}" 2>&1 | tail -3
mkdir -p gpurun_out; NFOPP_LANES=1 timeout 300 python scripts/tcs_trace.py > gpurun_out/tcs_trace.txt 2>&1; echo trace rc=$?
for lanes in 1 2; do
NFOPP_LANES=$lanes timeout 300 python bench.py --steps 100 --warmup 10 --no-extra --no-cpu-baseline --no-onf-microbench --no-solve > /tmp/b.json; python -c "
import json,sys
d=json.load(open('/tmp/b.json')); k=d['kernel_ms']
print('lanes$lanes value %.0f ms/step %.4f e2e %.0f | cand %.3f train %.3f path %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], k['onf_forward_candidates'], k['onf_train_fwd_bwd'], k['onf_path_fwd_dgrad']))"
done
