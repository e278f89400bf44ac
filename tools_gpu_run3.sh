#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone exit $?" > gpurun_out/summary3.txt; tail -3 gpurun_out/noalloc_alone.log >> gpurun_out/summary3.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_all.log 2>&1; echo "all exit $?" >> gpurun_out/summary3.txt; tail -3 gpurun_out/pytest_gpu_all.log >> gpurun_out/summary3.txt
cat gpurun_out/summary3.txt
