#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lstm_cluster.py -m gpu -q -x > gpurun_out/cluster.log 2>&1; echo "cluster exit $?"; tail -2 gpurun_out/cluster.log
for t in 1 2 3; do UMX_TEST_TERMS=$t timeout 900 python -m pytest tests/test_gpu_full.py -m gpu -q -x -s 2>&1 | grep -E "UMX-L chunk|passed|failed" ; done > gpurun_out/full_terms.log 2>&1; cat gpurun_out/full_terms.log
timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_default.log | cut -c1-400
timeout 1200 python bench.py --no-cpu-baseline --gemm-terms 1 > gpurun_out/bench_terms1.log 2>&1; tail -1 gpurun_out/bench_terms1.log | cut -c1-200
