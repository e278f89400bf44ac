#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lstm_cluster.py -m gpu -q -x 2>&1 | tail -2
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 2>&1 | tail -5
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 1 --seconds 50 --reps 1 2>&1 | tail -5
