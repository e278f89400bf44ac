#!/bin/bash
timeout 600 python tools/bench_target_parallel.py 1 420 2>&1 | tail -2
timeout 600 python tools/bench_target_parallel.py 2 420 2>&1 | tail -2
