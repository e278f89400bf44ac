#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2
timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
