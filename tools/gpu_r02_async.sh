#!/bin/bash
# two host-buffer calls in flight: parity test, then the bench step (e2e: one call at a time and pipelined)
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_properties.py tests/test_gpu_multiseg.py -m gpu -x -q > gpurun_out/async_pytest.log 2>&1; echo "pytest rc $?"; tail -5 gpurun_out/async_pytest.log
for i in 1 2; do
  UMX_BENCH_WATCHDOG_S=200 timeout -s KILL 300 python bench.py --no-cpu-baseline --steps 6 > gpurun_out/async_bench_$i.log 2> gpurun_out/async_bench_$i.err; echo "bench $i rc $?"
  tail -3 gpurun_out/async_bench_$i.err
done
