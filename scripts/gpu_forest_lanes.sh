#!/bin/bash
for lanes in 1 2 4; do
NFOPP_LANES=$lanes timeout 600 python bench.py --workload forest --problems 16384 --steps 20 --warmup 3 --no-extra --no-cpu-baseline --no-onf-microbench --no-solve > /tmp/f.json; python -c "
import json; d=json.load(open('/tmp/f.json')); print('forest16k lanes$lanes value %.0f ms/step %.3f e2e %.0f' % (d['value'], d['ms_per_step'], d['e2e']['value']))"
done
for lanes in 1 4; do
NFOPP_LANES=$lanes timeout 600 python bench.py --workload forest --steps 50 --warmup 5 --no-extra --no-cpu-baseline --no-onf-microbench --no-solve > /tmp/f.json; python -c "
import json; d=json.load(open('/tmp/f.json')); print('forest1k lanes$lanes value %.0f ms/step %.3f e2e %.0f' % (d['value'], d['ms_per_step'], d['e2e']['value']))"
done
