#!/bin/bash
# GPU call I: tf32 Gram kernel with TMA tensor-map staging: tests (both paths), config 4 timing A/B, SASS evidence
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tf32.py -x -q > gpurun_out/i_tf32_tests.log 2>&1
echo "tf32 tests (TMA where the table allows) rc=$?"; tail -5 gpurun_out/i_tf32_tests.log
DQ4ML_TF32_TMA=0 timeout 600 python -m pytest tests/test_gpu_tf32.py -x -q 2>&1 | tail -2
timeout 600 python tools/bench_configs.py --config 4 --gram-tier tf32x3 --reps 3 > gpurun_out/i_config4_tma.jsonl 2>&1
DQ4ML_TF32_TMA=0 timeout 600 python tools/bench_configs.py --config 4 --gram-tier tf32x3 --reps 3 > gpurun_out/i_config4_cpasync.jsonl 2>&1
cut -c1-420 gpurun_out/i_config4_tma.jsonl gpurun_out/i_config4_cpasync.jsonl
timeout 600 python -m pytest tests/test_gpu_weighted.py tests/test_gpu_huber.py -x -q > gpurun_out/i_weighted.log 2>&1
echo "weighted+huber rc=$?"; tail -5 gpurun_out/i_weighted.log
