#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/cluster.log 2>&1; echo "cluster exit $?" > gpurun_out/summary4.txt; tail -15 gpurun_out/cluster.log >> gpurun_out/summary4.txt
cat gpurun_out/summary4.txt
