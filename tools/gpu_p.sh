#!/bin/bash
# GPU call P: round-2 evidence: bench line, ncu launch list of the bench command, ncu --set full of the bench kernel and of the shipped filter kernel
set -x
mkdir -p gpurun_out
( timeout 600 python bench.py --no-cpu-baseline ) > gpurun_out/p_bench_n1.json 2> gpurun_out/p_bench_n1.err
cut -c1-300 gpurun_out/p_bench_n1.json
python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/p_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches.csv python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/p_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:stream_pipe_kernel -s 4 -c 2 -o gpurun_out/p_k2 python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/p_ncu2.log 2>&1
echo "k2 full rc=$?"
