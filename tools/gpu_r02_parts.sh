#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_multiseg.py -m gpu -q -x -k "not umxl" > gpurun_out/r02g_tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/r02g_tests.log
timeout 600 python bench.py --no-cpu-baseline --timeline gpurun_out/r02g_timeline.json > gpurun_out/r02g_bench.log 2>&1; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02g_bench.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, d['clocks']['sm_mhz'], 'd2h', d['e2e']['d2h_ms'])
r=d['roofline_in_order']; print('in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()})
PY
