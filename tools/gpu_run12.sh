#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/parity.log 2>&1; echo "parity exit $?"; tail -3 gpurun_out/parity.log
timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_default.log | cut -c1-300
timeout 1200 python bench.py --no-cpu-baseline --gemm-terms 1 > gpurun_out/bench_terms1.log 2>&1; tail -1 gpurun_out/bench_terms1.log | cut -c1-200
