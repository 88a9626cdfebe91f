#!/bin/bash
# GPU call D: in-kernel epilogue (host-mapped result, polling) + filter v3: tests, sweep, bench
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py tests/test_gpu_fit.py -x -q > gpurun_out/d_first_tests.log 2>&1
rc=$?
tail -15 gpurun_out/d_first_tests.log
[ $rc -ne 0 ] && exit 1
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -5 gpurun_out/d_pytest.log
timeout 600 python tools/bench_configs.py --config 5 --sizes 10000000,100000000,1000000000,4000000000 --reps 3 > gpurun_out/d_filter.jsonl 2>&1
cut -c1-260 gpurun_out/d_filter.jsonl
( time timeout 900 python bench.py ) > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err
echo "bench rc=$?"
tail -c 600 gpurun_out/d_bench.err
cut -c1-900 gpurun_out/d_bench.json
timeout 300 python tools/bench_configs.py --config 5 --sizes 1000000000 --reps 1 > gpurun_out/d_f5_plain.log 2>&1 && \
timeout 600 ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:dq_filter_stream -s 1 -c 1 -o gpurun_out/d_filter python tools/bench_configs.py --config 5 --sizes 1000000000 --reps 1 > gpurun_out/d_f5_ncu.log 2>&1
echo "ncu rc=$?"
