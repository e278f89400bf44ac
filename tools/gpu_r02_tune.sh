#!/bin/bash
mkdir -p gpurun_out
for fr in 32 48 64 96; do
  UMX_WI_FRAMES=$fr timeout -s KILL 300 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/r02p_bench_fr$fr.log 2>&1
  python - $fr <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/r02p_bench_fr{sys.argv[1]}.log').read().strip().splitlines()[-1])
r=d['roofline_in_order']; print('frames/CTA', sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'],1), 'in order', round(r['ms_per_step_in_order'],1), 'istft', round(r['stages']['istft']['ms_per_step'],1))
PY
done
