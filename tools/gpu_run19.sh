#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cli.py -m gpu -x -q -s > gpurun_out/cli.log 2>&1; echo "cli exit $?"; grep -E "target_|passed|failed|Error" gpurun_out/cli.log | tail -14
