#!/bin/bash
# 8-GPU call: the host copy ceiling at N = 8 and the track-sharded bench at N = 8
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_8gpu.txt 2>&1; nproc >> gpurun_out/r02_topo_8gpu.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 tools/host_copy_ceiling.py 2>/dev/null | tail -1 | tee gpurun_out/r02_ceiling_n8.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r02_bench_8gpu.json 2>gpurun_out/r02_bench_8gpu.err; echo "bench8 exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_8gpu.json').read().strip().splitlines()[-1])
print('N=8 value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],1), 'e2e ms', round(d['e2e']['ms_per_step'],1), 'h2d/d2h GB/s', d['e2e'].get('h2d_gbs'), d['e2e'].get('d2h_gbs'))
PY
