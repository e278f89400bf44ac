#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/parity.log 2>&1; echo "parity exit $?"; tail -3 gpurun_out/parity.log
for g in 4 2; do
UMX_LSTM_GROUPS=$g timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_g$g.log 2>&1; echo "bench exit $?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_g$g.log').read().strip().splitlines()[-1])
print('groups $g value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
UMX_LSTM_GROUPS=$g timeout 1200 python bench.py --no-cpu-baseline --pipeline 0 > gpurun_out/bench_g${g}_p0.log 2>&1; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_g${g}_p0.log').read().strip().splitlines()[-1])
print('groups $g pipeline 0 value', round(d['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
done
