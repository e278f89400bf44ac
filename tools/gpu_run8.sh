#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_step.py --tracks 1 --seconds 50 --reps 2"
timeout 600 $CMD > gpurun_out/prof_plain.log 2>&1 && timeout 1700 ncu --set full --clock-control none --import-source on --kernel-name regex:"lstm_cluster" -s 3 -c 1 -f -o gpurun_out/prof_lstm1 $CMD > gpurun_out/ncu_lstm.log 2>&1; echo "ncu lstm exit $?" > gpurun_out/summary8.txt
tail -2 gpurun_out/prof_plain.log >> gpurun_out/summary8.txt
cat gpurun_out/summary8.txt
