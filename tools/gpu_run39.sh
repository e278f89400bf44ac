#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/bench_configs.py > gpurun_out/configs.log 2>&1; echo "exit $?"; cat gpurun_out/configs.log | tail -12
timeout 600 python tools/prof_step.py --tracks 1 --seconds 420 --reps 2 2>&1 | tail -2
