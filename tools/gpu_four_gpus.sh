#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 4 --steps 2 --warmup 3 > gpurun_out/bench_4gpu.log 2>gpurun_out/bench_4gpu.err; echo "bench4 exit $?"; tail -n 1 gpurun_out/bench_4gpu.log | cut -c1-260
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 4 --steps 1 --warmup 0 2>/dev/null | tail -n 1 | cut -c1-200
