#!/bin/bash
# GPU call F: filter v5 (publish early, look back late, one barrier), full tests, bench
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q > gpurun_out/f_filter_tests.log 2>&1
rc=$?
tail -4 gpurun_out/f_filter_tests.log
if [ $rc -eq 0 ]; then
  for shape in 8 6; do
    DQ4ML_FILTER_SHAPE=$shape timeout 300 python -m pytest tests/test_gpu_dq.py -x -q 2>&1 | tail -1
    DQ4ML_FILTER_SHAPE=$shape timeout 600 python tools/bench_configs.py --config 5 --sizes 10000000,100000000,1000000000,4000000000 --reps 3 > gpurun_out/f_filter_shape$shape.jsonl 2>&1
    cut -c1-200 gpurun_out/f_filter_shape$shape.jsonl
  done
fi
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/f_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -5 gpurun_out/f_pytest.log
( time timeout 900 python bench.py --no-extra --no-cpu-baseline ) > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
cut -c1-1500 gpurun_out/f_bench.json
timeout 300 python tools/bench_configs.py --config 5 --sizes 1000000000 --reps 1 > gpurun_out/f_f5_plain.log 2>&1 && \
timeout 600 ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:dq_filter_stream -s 1 -c 1 -o gpurun_out/f_filter python tools/bench_configs.py --config 5 --sizes 1000000000 --reps 1 > gpurun_out/f_f5_ncu.log 2>&1
echo "ncu rc=$?"
