#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/prof_step.py --tracks 16 --seconds 50 --reps 1"
timeout 600 $CMD > gpurun_out/prof_plain.log 2>&1 && timeout 1700 ncu --set full --clock-control none --import-source on --kernel-name regex:"wiener_reduce|wiener_apply|istft_kernel|stft_kernel" -c 4 -f -o gpurun_out/prof_hbm $CMD > gpurun_out/ncu_hbm.log 2>&1; echo "ncu exit $?"
