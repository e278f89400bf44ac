#!/bin/bash
mkdir -p gpurun_out
python -c "
import __graft_entry__ as g
u=g.load_package(); print('max active lstm clusters:', u.lib().umx_cuda_lstm_max_active_clusters())"
timeout 1200 python bench.py --no-cpu-baseline --tracks 16 > gpurun_out/bench_t16.log 2>&1; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_t16.log').read().strip().splitlines()[-1])
print('t16 value', round(d['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
