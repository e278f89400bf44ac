#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x -s > gpurun_out/all_gpu.log 2>&1; echo "all gpu exit $?"; grep -E "UMX-L chunk|cli target|dropin target|passed|failed|skipped" gpurun_out/all_gpu.log | tail -20
