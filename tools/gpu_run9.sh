#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/cluster.log 2>&1; echo "cluster exit $?"; tail -3 gpurun_out/cluster.log
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 1 --seconds 50 --reps 1 > gpurun_out/dbg1.log 2>&1
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 16 --seconds 50 --reps 1 > gpurun_out/dbg16.log 2>&1
tail -2 gpurun_out/dbg1.log; tail -2 gpurun_out/dbg16.log
