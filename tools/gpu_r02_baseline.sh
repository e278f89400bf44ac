#!/bin/bash
# round 2 first call: every GPU test (incl. the multi-segment comparison with the reference CLI) + a default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/r02_all_gpu.log 2>&1; echo "all gpu exit $?"
grep -E "multiseg|overlap \[|passed|failed|skipped|Error" gpurun_out/r02_all_gpu.log | tail -20
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02_bench_start.log 2>&1; echo "bench exit $?"; tail -c 3000 gpurun_out/r02_bench_start.log
