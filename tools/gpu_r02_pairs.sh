#!/bin/bash
# CTA-pair GEMM: parity, then in-order and wavefront bench lines against the single-CTA kernel
mkdir -p gpurun_out
UMX_TEST_PAIRS=1 timeout -s KILL 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "golden or target_mags or multi_segment or cta_pairs" > gpurun_out/r02n_pair_tests.log 2>&1; echo "pair tests exit $?"; tail -6 gpurun_out/r02n_pair_tests.log
for impl in 0 2; do
  timeout -s KILL 300 python bench.py --no-cpu-baseline --no-e2e --gemm-impl $impl > gpurun_out/r02n_bench_impl$impl.log 2>&1; echo "bench impl $impl exit $?"
  python - $impl <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r02n_bench_impl{sys.argv[1]}.log').read().strip().splitlines()[-1])
    print('impl', sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'],1)); r=d['roofline_in_order']; print('   in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()})
except Exception as e:
    print('failed', e); print(open(f'gpurun_out/r02n_bench_impl{sys.argv[1]}.log').read()[-800:])
PY
done
