#!/bin/bash
# GPU call G: filter v6 variant A/B (v1 structure + switchable changes), huber tests, full tests, bench
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q > gpurun_out/g_filter_tests.log 2>&1
rc=$?
tail -4 gpurun_out/g_filter_tests.log
: > gpurun_out/g_filter_variants.jsonl
if [ $rc -eq 0 ]; then
  for v in 0 1 2 4 3 5 6 7; do
    echo "{\"variant\": $v}" >> gpurun_out/g_filter_variants.jsonl
    DQ4ML_FILTER_VARIANT=$v timeout 120 python -m pytest tests/test_gpu_dq.py -x -q 2>&1 | tail -1
    DQ4ML_FILTER_VARIANT=$v timeout 300 python tools/bench_configs.py --config 5 --sizes 100000000,1000000000 --reps 3 >> gpurun_out/g_filter_variants.jsonl 2>&1
  done
  for v in 0 7; do
    echo "{\"variant\": $v, \"shape\": 6}" >> gpurun_out/g_filter_variants.jsonl
    DQ4ML_FILTER_SHAPE=6 DQ4ML_FILTER_VARIANT=$v timeout 300 python tools/bench_configs.py --config 5 --sizes 100000000,1000000000 --reps 3 >> gpurun_out/g_filter_variants.jsonl 2>&1
  done
  cut -c1-130 gpurun_out/g_filter_variants.jsonl
fi
timeout 600 python -m pytest tests/test_gpu_huber.py -x -q > gpurun_out/g_huber.log 2>&1; tail -30 gpurun_out/g_huber.log
( time timeout 900 python -m pytest tests -m gpu -q ) > gpurun_out/g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/g_pytest.log
tail -8 gpurun_out/g_pytest.log
