#!/bin/bash
mkdir -p gpurun_out
for lt in 4 2 1; do
  timeout 600 python bench.py --no-cpu-baseline --no-in-order --lane-targets $lt --timeline gpurun_out/r02f_timeline_l$lt.json > gpurun_out/r02f_bench_l$lt.log 2>&1; echo "bench l$lt exit $?"
  python - "$lt" <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/r02f_bench_l{sys.argv[1]}.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, d['clocks']['sm_mhz'])
PY
done
timeout 300 python bench.py --no-cpu-baseline --no-in-order --tracks 1 --track-seconds 420 --steps 5 > gpurun_out/r02f_single7.log 2>&1; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02f_single7.log').read().strip().splitlines()[-1])
print('single 7 min', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
