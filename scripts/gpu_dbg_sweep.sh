#!/bin/bash
# bench lines of the corridor workload under the NFOPP_TCS_DBG experiment switches of onf_tcs.cu
for dbg in ${DBGS:-0 4 8 12}; do
for lanes in ${LANES:-1 2}; do
NFOPP_TCS_DBG=$dbg NFOPP_LANES=$lanes timeout 300 python bench.py --steps 100 --warmup 10 --no-extra --no-cpu-baseline --no-onf-microbench --no-solve > /tmp/b.json; python -c "
import json,sys
d=json.load(open('/tmp/b.json')); k=d['kernel_ms']
print('dbg$dbg lanes$lanes value %.0f ms/step %.4f e2e %.0f | cand %.3f train %.3f path %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], k['onf_forward_candidates'], k['onf_train_fwd_bwd'], k['onf_path_fwd_dgrad']))"
done
done
