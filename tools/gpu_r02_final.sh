#!/bin/bash
# end-of-round evidence on one GPU: every GPU test, smoke, final bench lines (default and gemm_terms 2), the reference arm, other
# configs, the ncu launch list of one bench step
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x -s > gpurun_out/r02_final_tests.log 2>&1; echo "all gpu tests exit $?"; grep -E "passed|failed|skipped" gpurun_out/r02_final_tests.log | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_final_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r02_final_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2>gpurun_out/r02_bench_1gpu.err; echo "bench exit $?"
timeout 600 python bench.py --steps 10 --warmup 3 --gemm-terms 2 --no-cpu-baseline > gpurun_out/r02_bench_1gpu_terms2.json 2>/dev/null; echo "bench terms2 exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null; echo "reference arm exit $?"
for cfg in "1 10" "1 420"; do set -- $cfg; timeout 300 python bench.py --no-cpu-baseline --no-in-order --tracks $1 --track-seconds $2 --steps 5 > gpurun_out/r02_bench_tracks$1_$2s.json 2>/dev/null; done
python - <<'PY'
import json
for f in ('r02_bench_1gpu','r02_bench_1gpu_terms2','r02_bench_tracks1_10s','r02_bench_tracks1_420s'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    print(f, 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],2), 'e2e ms', round(d['e2e']['ms_per_step'],2), d['clocks'].get('sm_mhz'))
    r=d.get('roofline_in_order')
    if r: print('   in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()}, 'issued', round(r['frac_issued'],3))
    if 'cpu_baseline' in d: print('   cpu', d['cpu_baseline']['value'])
d=json.loads(open('gpurun_out/r02_bench_reference_arm.json').read().strip().splitlines()[-1]); print('reference arm', d['value'], d.get('stock_window_check'))
PY
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_bench_step.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-in-order --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
python tools/bench_cli.py 180 420 > gpurun_out/r02_cli.txt 2>&1; cat gpurun_out/r02_cli.txt
