#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lstm_cluster.py tests/test_gpu_options.py tests/test_gpu_model_load.py -m gpu -q -x > gpurun_out/r02o_tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/r02o_tests.log
for impl in 0 2 3; do
  timeout -s KILL 300 python bench.py --no-cpu-baseline --gemm-impl $impl > gpurun_out/r02o_bench_impl$impl.log 2>&1; echo "bench impl $impl exit $?"
  python - $impl <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/r02o_bench_impl{sys.argv[1]}.log').read().strip().splitlines()[-1])
print('impl', sys.argv[1], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1)); r=d['roofline_in_order']; print('   in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()}, 'issued frac', round(r['frac_issued'],3))
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02_gemm_wih_pair python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm exit $?"
