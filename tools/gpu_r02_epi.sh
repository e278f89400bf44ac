#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full.py tests/test_gpu_cli.py -m gpu -q -x > gpurun_out/r02l_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r02l_tests.log
timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/r02l_bench.log 2>&1; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02l_bench.log').read().strip().splitlines()[-1])
print('value', round(d['value']), 'ms', round(d['ms_per_step'],1)); r=d['roofline_in_order']; print('in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()})
PY
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_bench_step.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-in-order --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02_gemm_wih python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm exit $?"
python tools/bench_cli.py 180 > gpurun_out/r02_cli.txt 2>&1; cat gpurun_out/r02_cli.txt
