#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/prof_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 0 -c 1 -f -o gpurun_out/gemm3_fc1_r01 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_a.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:wiener_apply -c 1 -f -o gpurun_out/wiener_apply_r01 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_b.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:istft_kernel -c 1 -f -o gpurun_out/istft_final_r01 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_c.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:stft_kernel -c 1 -f -o gpurun_out/stft_r01 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_d.log 2>&1
echo "exit $?"; ls -la gpurun_out/*_r01.ncu-rep | tail -6
