#!/bin/bash
mkdir -p gpurun_out
UMX_NVCC_EXTRA="-DUMX_LSTM_TIMING" python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/build_t.log 2>&1; echo "build exit $?"
UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 2 2>&1 | grep -E "cycles/step|launches" | tail -4
