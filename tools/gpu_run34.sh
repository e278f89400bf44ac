#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipeline or multi_segment" 2>&1 | tail -3
for pl in 0 2; do
timeout 1200 python bench.py --no-cpu-baseline --pipeline $pl > gpurun_out/bench_pl$pl.log 2>&1; echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_pl$pl.log').read().strip().splitlines()[-1])
print('pipeline $pl value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, d['clocks']['sm_mhz'])
PY
done
