#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_target_parallel.py -m gpu -q -x 2>&1 | tail -3
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_2gpu.log 2>gpurun_out/bench_2gpu.err; echo "bench2 exit $?"; tail -n 1 gpurun_out/bench_2gpu.log | cut -c1-400
