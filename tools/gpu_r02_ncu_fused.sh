#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/prof_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:wiener_istft_kernel -c 1 -f -o gpurun_out/r02_wiener_istft python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_wi.log 2>&1
echo "exit $?"; tail -2 gpurun_out/ncu_wi.log
