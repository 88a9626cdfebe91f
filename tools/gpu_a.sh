#!/bin/bash
# round-2 GPU call A: full GPU test suite, default bench (with extra configs), reference arm, filter stall profile
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
nproc >> gpurun_out/a_gpu.txt
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
( time python bench.py ) > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench rc=$?"
tail -c 3000 gpurun_out/a_bench.err
( time python bench.py --impl reference --steps 5 --warmup 1 ) > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err
python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/a_f5_plain.log 2>&1 && \
ncu --section SpeedOfLight --section WarpStateStats --section SourceCounters --section Occupancy --section MemoryWorkloadAnalysis --clock-control none --import-source on -k regex:dq_filter_compact -s 1 -c 1 -o gpurun_out/a_filter python tools/bench_configs.py --config 5 --sizes 100000000 --reps 1 > gpurun_out/a_f5_ncu.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out
