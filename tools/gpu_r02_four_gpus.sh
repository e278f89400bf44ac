#!/bin/bash
# 4-GPU call: target-parallel parity (4 stems on 4 GPUs), its timing (1 track and an 8-track batch), the host copy ceiling at
# N = 1, 2, 4 and the sharded bench at N = 4
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_4gpu.txt 2>&1; nproc >> gpurun_out/r02_topo_4gpu.txt
timeout 600 python -m pytest tests/test_gpu_target_parallel.py -m gpu -q -x > gpurun_out/r02_tp_test.log 2>&1; echo "tp test exit $?"; tail -2 gpurun_out/r02_tp_test.log
for w in 1 2 4; do timeout 300 python tools/bench_target_parallel.py $w 420 1 2>&1 | grep target-parallel | tee -a gpurun_out/r02_tp_bench.log; done
for w in 1 4; do timeout 300 python tools/bench_target_parallel.py $w 420 8 2>&1 | grep target-parallel | tee -a gpurun_out/r02_tp_bench.log; done
timeout 120 python tools/host_copy_ceiling.py | tee gpurun_out/r02_ceiling_n1.json
for n in 2 4; do timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/host_copy_ceiling.py 2>/dev/null | tail -1 | tee gpurun_out/r02_ceiling_n$n.json; done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/r02_bench_4gpu.json 2>gpurun_out/r02_bench_4gpu.err; echo "bench4 exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_4gpu.json').read().strip().splitlines()[-1])
print('N=4 value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), 'h2d/d2h GB/s', d['e2e'].get('h2d_gbs'), d['e2e'].get('d2h_gbs'))
PY
