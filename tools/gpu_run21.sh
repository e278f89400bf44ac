#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_target_parallel.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/tp.log 2>&1; echo "tp exit $?"; tail -15 gpurun_out/tp.log
