#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/bench_s1.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01_launches.csv python bench.py --no-cpu-baseline --steps 1 --warmup 1 > gpurun_out/ncu_launches.log 2>&1
echo "exit $?"; tail -1 gpurun_out/bench_s1.log | cut -c1-300
