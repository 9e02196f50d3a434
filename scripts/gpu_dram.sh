#!/bin/bash
# DRAM traffic and duration of the three streamed ONF launches of a step (ncu, one lane), then bench lines
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-onf-microbench --no-solve --no-extra"
mkdir -p gpurun_out
NFOPP_LANES=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:onf_tcs_kernel -s 9 -c 3 --csv --log-file gpurun_out/dram.csv $B > gpurun_out/dram.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/dram.csv')) if r and not r[0].startswith('==')]
h=rows[0]; 
for r in rows[1:]:
    print(r[h.index('ID')], r[h.index('Metric Name')], r[h.index('Metric Value')], r[h.index('Metric Unit')])
PY
LANES="${LANES:-1 4}" DBGS=0 scripts/gpu_dbg_sweep.sh
