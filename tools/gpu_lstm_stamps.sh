#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/parity.log 2>&1; echo "parity exit $?"; tail -3 gpurun_out/parity.log
timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step']), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
UMX_NVCC_EXTRA="-DUMX_LSTM_TIMING" python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/build_t.log 2>&1; echo "build exit $?"
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 2 2>&1 | grep -E -A4 "lstm_cluster<16>|launches" | tail -6
