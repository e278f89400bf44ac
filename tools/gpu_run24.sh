#!/bin/bash
mkdir -p gpurun_out
for eg in 1 2; do
  UMX_NVCC_EXTRA="-DUMX_LSTM_TIMING -DUMX_LSTM_EPI_GROUPS=$eg" python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/build_$eg.log 2>&1; echo "build $eg exit $?"
  UMX_LSTM_DEBUG=1 timeout 600 python tools/prof_step.py --tracks 32 --seconds 50 --reps 1 2>&1 | grep -A3 "lstm_cluster<16>" | head -8
done
