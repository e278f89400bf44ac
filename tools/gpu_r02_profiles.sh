#!/bin/bash
# round-2 evidence: CLI / drop-in wall clock, final bench lines (default, gemm_terms 2), ncu launch list of one bench step,
# ncu --set full of the W_ih GEMM and the recurrence
mkdir -p gpurun_out
python tools/bench_cli.py 180 420 > gpurun_out/r02_cli.txt 2>&1; cat gpurun_out/r02_cli.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2>gpurun_out/r02_bench_1gpu.err; echo "bench exit $?"
timeout 600 python bench.py --steps 5 --warmup 3 --gemm-terms 2 --no-cpu-baseline > gpurun_out/r02_bench_1gpu_terms2.json 2>/dev/null; echo "bench terms2 exit $?"
python - <<'PY'
import json
for f in ('r02_bench_1gpu','r02_bench_1gpu_terms2'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    print(f, 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), d['clocks'])
    r=d['roofline_in_order']; print('   in order', round(r['ms_per_step_in_order'],1), {k: (round(v['ms_per_step'],1), round(v['frac'],3)) for k,v in r['stages'].items()})
    if 'cpu_baseline' in d: print('   cpu', d['cpu_baseline']['value'])
PY
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 1100 -c 500 --csv --log-file gpurun_out/r02_launches_bench_step.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-in-order --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02_gemm_wih python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lstm_cluster_kernel -c 1 -f -o gpurun_out/r02_lstm_cluster python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_lstm.log 2>&1; echo "ncu lstm exit $?"
