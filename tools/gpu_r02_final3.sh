#!/bin/bash
# last check of the round: GPU tests, smoke, the default bench command twice (as the driver runs it)
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/smoke.log
UMX_BENCH_WATCHDOG_S=400 timeout -s KILL 500 python bench.py > gpurun_out/final_default.log 2> gpurun_out/final_default.err; echo "bench (default flags) rc $?"
for i in 1 2; do
  UMX_BENCH_WATCHDOG_S=150 timeout -s KILL 240 python bench.py --no-cpu-baseline --steps 6 > gpurun_out/final3_$i.log 2> gpurun_out/final3_$i.err; echo "bench $i rc $?"
done
