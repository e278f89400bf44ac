#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" > gpurun_out/summary5.txt; tail -4 gpurun_out/pytest_gpu.log >> gpurun_out/summary5.txt
timeout 1200 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench exit $?" >> gpurun_out/summary5.txt; tail -1 gpurun_out/bench_default.log | cut -c1-1500 >> gpurun_out/summary5.txt
timeout 900 python bench.py --tracks 1 --track-seconds 420 --no-cpu-baseline > gpurun_out/bench_single7min.log 2>&1; echo "bench1 exit $?" >> gpurun_out/summary5.txt; tail -1 gpurun_out/bench_single7min.log | cut -c1-1200 >> gpurun_out/summary5.txt
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --tracks 16"
timeout 900 $CMD > gpurun_out/plain_for_ncu.log 2>&1 && timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"stft|gemm_tc|lstm|wiener|istft|finalize" -c 300 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_run.log 2>&1; echo "ncu exit $?" >> gpurun_out/summary5.txt
cat gpurun_out/summary5.txt
