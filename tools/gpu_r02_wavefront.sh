#!/bin/bash
# wavefront schedule: bit-identity with the in-order schedule, then bench lines per schedule
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lstm_cluster.py tests/test_gpu_properties.py tests/test_gpu_model_load.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_wf_tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/r02_wf_tests.log
for cfg in "3 1" "3 2" "3 4" "2 1"; do
  set -- $cfg
  timeout 600 python bench.py --no-cpu-baseline --no-in-order --pipeline $1 --lane-targets $2 > gpurun_out/r02_bench_p$1_l$2.log 2>&1; echo "bench p$1 l$2 exit $?"
  python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r02_bench_p{sys.argv[1]}_l{sys.argv[2]}.log').read().strip().splitlines()[-1])
    print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, 'h2d/d2h ms', d['e2e'].get('h2d_ms'), d['e2e'].get('d2h_ms'), d['clocks'])
except Exception as e:
    print('parse failed', e); print(open(f'gpurun_out/r02_bench_p{sys.argv[1]}_l{sys.argv[2]}.log').read()[-1500:])
PY
done
python tools/bench_configs.py > gpurun_out/r02_other_configs.log 2>&1; cat gpurun_out/r02_other_configs.log | tail -8
