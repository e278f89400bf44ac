#!/bin/bash
# UMX-L full-size parity against the compiled reference for each gemm_terms setting
mkdir -p gpurun_out
: > gpurun_out/parity_by_terms.log
for t in 3 2 1; do
  echo "gemm_terms=$t" >> gpurun_out/parity_by_terms.log
  UMX_TEST_TERMS=$t timeout 900 python -m pytest tests/test_gpu_full.py -m gpu -q -x -s 2>&1 | grep -E "UMX-L chunk|passed|failed" >> gpurun_out/parity_by_terms.log
done
cat gpurun_out/parity_by_terms.log
