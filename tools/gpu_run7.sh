#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/cluster.log 2>&1; echo "cluster exit $?" > gpurun_out/summary7.txt; tail -12 gpurun_out/cluster.log >> gpurun_out/summary7.txt
timeout 600 python tools/prof_step.py --tracks 16 --seconds 50 --reps 2 > gpurun_out/prof16.log 2>&1; tail -1 gpurun_out/prof16.log >> gpurun_out/summary7.txt
timeout 600 python tools/prof_step.py --tracks 1 --seconds 50 --reps 2 > gpurun_out/prof1.log 2>&1; tail -1 gpurun_out/prof1.log >> gpurun_out/summary7.txt
cat gpurun_out/summary7.txt
