#!/bin/bash
# GPU call C: filter kernel v2 (direct stores): tests, sweep, ncu
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q > gpurun_out/c_filter_tests.log 2>&1
rc=$?
tail -4 gpurun_out/c_filter_tests.log
[ $rc -ne 0 ] && exit 1
timeout 600 python tools/bench_configs.py --config 5 --sizes 10000000,100000000,1000000000,4000000000 --reps 3 > gpurun_out/c_filter.jsonl 2>&1
cut -c1-260 gpurun_out/c_filter.jsonl
timeout 300 python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/c_f5_plain.log 2>&1 && \
timeout 600 ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:dq_filter_stream -s 1 -c 1 -o gpurun_out/c_filter python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/c_f5_ncu.log 2>&1
echo "ncu rc=$?"
