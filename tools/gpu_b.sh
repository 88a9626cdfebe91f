#!/bin/bash
# round-2 GPU call B: new persistent filter kernel (tests first, under a short timeout), bench, A/B of the two filter kernels, ncu
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q > gpurun_out/b_filter_tests.log 2>&1
rc=$?
echo "filter tests rc=$rc" >> gpurun_out/b_filter_tests.log
tail -15 gpurun_out/b_filter_tests.log
if [ $rc -ne 0 ]; then
  echo "new filter kernel failed its tests: re-running with the legacy kernel to isolate"
  DQ4ML_FILTER_LEGACY=1 timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5
  exit 1
fi
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -5 gpurun_out/b_pytest.log
timeout 600 python tools/bench_configs.py --config 5 --sizes 10000000,100000000,1000000000,4000000000 --reps 3 > gpurun_out/b_filter_new.jsonl 2>&1
DQ4ML_FILTER_LEGACY=1 timeout 600 python tools/bench_configs.py --config 5 --sizes 100000000,1000000000 --reps 3 > gpurun_out/b_filter_legacy.jsonl 2>&1
cat gpurun_out/b_filter_new.jsonl gpurun_out/b_filter_legacy.jsonl | cut -c1-330
( time timeout 900 python bench.py ) > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err
echo "bench rc=$?"
tail -c 1500 gpurun_out/b_bench.err
timeout 300 python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/b_f5_plain.log 2>&1 && \
timeout 600 ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:dq_filter_stream -s 1 -c 1 -o gpurun_out/b_filter python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/b_f5_ncu.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out | head -40
