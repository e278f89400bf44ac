#!/bin/bash
# the default bench step (device loop, e2e one call at a time and two in flight, in-order leg) over and over
mkdir -p gpurun_out
ok=0; bad=0
for i in $(seq 1 ${1:-7}); do
  UMX_BENCH_WATCHDOG_S=150 timeout -s KILL 240 python bench.py --no-cpu-baseline --steps 5 > gpurun_out/soak2_$i.log 2> gpurun_out/soak2_$i.err; rc=$?
  if [ $rc -eq 0 ]; then ok=$((ok+1)); else bad=$((bad+1)); echo "run $i rc $rc"; grep -E "Timeout|File \"/root/repo" gpurun_out/soak2_$i.err | head -6; fi
done
echo "default bench: $ok ok, $bad failed"
