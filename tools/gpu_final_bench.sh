#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python bench.py > gpurun_out/bench_final.log 2>gpurun_out/bench_final.err; echo "bench exit $?"; tail -c 400 gpurun_out/bench_final.log
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>gpurun_out/bench_ref.err; echo "ref exit $?"
