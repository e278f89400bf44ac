#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_full.py -m gpu -q -x > gpurun_out/r02k_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r02k_tests.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02k_bench.log 2>&1; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02k_bench.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, d['clocks']['sm_mhz'])
r=d['roofline_in_order']; print('in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()})
PY
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:wiener_istft_kernel -s 1 -c 1 -f -o gpurun_out/r02_wiener_istft python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_wi.log 2>&1
echo "ncu exit $?"
