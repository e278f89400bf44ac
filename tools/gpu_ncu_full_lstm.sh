#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/prof_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:lstm_cluster -s 1 -c 1 -f -o gpurun_out/lstm_r01b python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_lstm.log 2>&1
echo "exit $?"; tail -3 gpurun_out/ncu_lstm.log
