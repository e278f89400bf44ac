#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline --tracks 1 --track-seconds 420 > gpurun_out/bench_cfg3.log 2>&1; echo "exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_cfg3.log').read().strip().splitlines()[-1])
print('config 3 (1 x 7 min): value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
timeout 900 python bench.py --no-cpu-baseline --tracks 1 --track-seconds 10 > gpurun_out/bench_cfg1.log 2>&1; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_cfg1.log').read().strip().splitlines()[-1])
print('config 1 (1 x 10 s): value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
