#!/bin/bash
# bring-up run: each group in its own process so a trapped kernel cannot poison the rest
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
T=tests/test_gpu_parity.py
run() { name=$1; shift; timeout 900 python -m pytest -q -m gpu "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt; tail -3 gpurun_out/$name.log >> gpurun_out/summary.txt; }
run base "$T" -k "not inference and not target_mags and not demix"
run simt "$T::test_umx_inference_golden[1]" "$T::test_target_mags_vs_oracle[1]" "$T::test_demix_multi_segment_multi_track[1-3]" "$T::test_demix_shift_offsets_and_trim"
run tc "$T::test_umx_inference_golden[0]" "$T::test_target_mags_vs_oracle[0]" "$T::test_demix_multi_segment_multi_track[0-3]" "$T::test_demix_multi_segment_multi_track[0-2]" "$T::test_demix_multi_segment_multi_track[0-1]"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
