#!/bin/bash
# per-kernel times of the small kernels (one lane) + the 4-lane step
for i in 1 2; do LANES="1" DBGS=0 scripts/gpu_dbg_sweep.sh > /dev/null; python -c "
import json; k=json.load(open('/tmp/b.json'))['kernel_ms']; print('resample %.4f sample_pre %.4f traj_post %.4f traj_pre %.4f tick %.4f' % (k['resample_label'], k['sample_pre'], k['traj_post'], k['traj_pre'], k['tick']))"; done
LANES=4 DBGS=0 scripts/gpu_dbg_sweep.sh
