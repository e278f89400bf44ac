#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/prof_plain.log 2>&1 && \
timeout 1500 ncu --clock-control none -k regex:gemm_tc --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,launch__grid_size --csv --log-file gpurun_out/gemm_metrics.csv python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 > gpurun_out/ncu_gemm.log 2>&1
echo "exit $?"; tail -2 gpurun_out/ncu_gemm.log
