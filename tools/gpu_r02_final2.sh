#!/bin/bash
# final check: GPU tests, smoke, then the default bench step (device loop, e2e loop, in-order leg) over and over
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/smoke.log
ok=0; bad=0
for i in $(seq 1 ${1:-10}); do
  UMX_BENCH_WATCHDOG_S=150 timeout -s KILL 240 python bench.py --no-cpu-baseline --steps 6 > gpurun_out/final_$i.log 2> gpurun_out/final_$i.err; rc=$?
  if [ $rc -eq 0 ]; then ok=$((ok+1)); else bad=$((bad+1)); echo "run $i rc $rc"; grep -E "Timeout|File \"/root/repo" gpurun_out/final_$i.err | head -6; fi
done
echo "default bench: $ok ok, $bad failed"
