#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" > gpurun_out/summary.txt; tail -5 gpurun_out/pytest_gpu.log >> gpurun_out/summary.txt
timeout 900 python bench.py --tracks 8 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_t8.log 2>&1; echo "bench8 exit $?" >> gpurun_out/summary.txt; tail -2 gpurun_out/bench_t8.log >> gpurun_out/summary.txt
timeout 1200 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt; tail -2 gpurun_out/bench_default.log >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "benchref exit $?" >> gpurun_out/summary.txt; tail -1 gpurun_out/bench_ref.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
