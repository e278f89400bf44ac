#!/bin/bash
# odd batch sizes and a longer run through the default schedule: no traps, no hangs, sane throughput
mkdir -p gpurun_out
for cfg in "17 100" "33 70" "5 240"; do
  set -- $cfg
  timeout 900 python bench.py --no-cpu-baseline --tracks $1 --track-seconds $2 --steps 2 --warmup 1 > gpurun_out/soak_$1.log 2>&1; rc=$?
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/soak_$1.log').read().strip().splitlines()[-1])
    print('tracks $1 x $2 s: exit $rc value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1))
except Exception as e:
    print('tracks $1: exit $rc FAILED', e); print(open('gpurun_out/soak_$1.log').read()[-600:])
PY
done
timeout 1200 python bench.py --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/soak_long.log 2>&1; rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/soak_long.log').read().strip().splitlines()[-1])
print('32 x 180 s, 10 steps: exit $rc value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), d['clocks'])
PY
