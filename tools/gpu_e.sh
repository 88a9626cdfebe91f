#!/bin/bash
# GPU call E: filter v4 shapes, epilogue A/B (host path breakdown), full tests, bench
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dq.py tests/test_zz_no_alloc.py -x -q > gpurun_out/e_filter_tests.log 2>&1
rc=$?
tail -4 gpurun_out/e_filter_tests.log
if [ $rc -eq 0 ]; then
  for shape in 8 6 4; do
    DQ4ML_FILTER_SHAPE=$shape timeout 300 python -m pytest tests/test_gpu_dq.py -x -q 2>&1 | tail -1
    DQ4ML_FILTER_SHAPE=$shape timeout 600 python tools/bench_configs.py --config 5 --sizes 100000000,1000000000,4000000000 --reps 3 > gpurun_out/e_filter_shape$shape.jsonl 2>&1
    cut -c1-200 gpurun_out/e_filter_shape$shape.jsonl
  done
fi
timeout 300 python tools/step_overhead.py > gpurun_out/e_overhead_on.json 2>&1; cat gpurun_out/e_overhead_on.json
DQ4ML_EPILOGUE=off timeout 300 python tools/step_overhead.py > gpurun_out/e_overhead_off.json 2>&1; cat gpurun_out/e_overhead_off.json
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/e_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/e_pytest.log
tail -5 gpurun_out/e_pytest.log
( time timeout 900 python bench.py --no-extra --no-cpu-baseline ) > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err
cut -c1-1200 gpurun_out/e_bench.json
