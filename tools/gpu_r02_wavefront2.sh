#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipeline_modes" > gpurun_out/r02_wf_tests2.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/r02_wf_tests2.log
for cfg in "3 1" "3 2" "3 4" "2 1"; do
  set -- $cfg
  timeout 600 python bench.py --no-cpu-baseline --no-in-order --pipeline $1 --lane-targets $2 > gpurun_out/r02b_bench_p$1_l$2.log 2>&1; echo "bench p$1 l$2 exit $?"
  python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r02b_bench_p{sys.argv[1]}_l{sys.argv[2]}.log').read().strip().splitlines()[-1])
    print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()}, 'h2d/d2h ms', d['e2e'].get('h2d_ms'), d['e2e'].get('d2h_ms'), d['clocks']['sm_mhz'])
except Exception as e:
    print('parse failed', e); print(open(f'gpurun_out/r02b_bench_p{sys.argv[1]}_l{sys.argv[2]}.log').read()[-1500:])
PY
done
for lt in 1 4; do
timeout 300 python bench.py --no-cpu-baseline --no-in-order --tracks 1 --track-seconds 420 --lane-targets $lt --steps 5 > gpurun_out/r02b_bench_single7min_l$lt.log 2>&1; python - $lt <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/r02b_bench_single7min_l{sys.argv[1]}.log').read().strip().splitlines()[-1])
print('single 7 min lane', sys.argv[1], 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), {k: round(v['ms_per_step'],1) for k,v in d['stages'].items()})
PY
done
python tools/host_copy_ceiling.py > gpurun_out/r02b_ceiling_1gpu_2d.json; cat gpurun_out/r02b_ceiling_1gpu_2d.json
python tools/host_copy_ceiling.py --d2h-1d > gpurun_out/r02b_ceiling_1gpu_1d.json; cat gpurun_out/r02b_ceiling_1gpu_1d.json
python tools/host_copy_ceiling.py --one-direction d2h > gpurun_out/r02b_ceiling_1gpu_d2h.json; cat gpurun_out/r02b_ceiling_1gpu_d2h.json
nvidia-smi topo -m > gpurun_out/r02b_topo.txt 2>&1; lscpu | head -30 >> gpurun_out/r02b_topo.txt; numactl -H >> gpurun_out/r02b_topo.txt 2>&1; tail -40 gpurun_out/r02b_topo.txt
