#!/bin/bash
mkdir -p gpurun_out
for T in 3 1; do
CMD="python tools/prof_step.py --tracks 16 --seconds 50 --reps 1 --terms $T"
timeout 600 $CMD > gpurun_out/prof_plain_t$T.log 2>&1 && timeout 1700 ncu --set full --clock-control none --import-source on --kernel-name regex:"gemm_tc" -s 1 -c 1 -f -o gpurun_out/prof_gemm_t$T $CMD > gpurun_out/ncu_gemm_t$T.log 2>&1; echo "ncu gemm t$T exit $?"
done
