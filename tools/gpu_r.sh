#!/bin/bash
# GPU call R: ncu full capture (+ warp states, source counters) of the producer-warp fp64 Gram kernel at 8 M x 256
set -x
mkdir -p gpurun_out
python tools/bench_configs.py --config 4 --rows 8000000 --reps 1 --gram-tier fp64 2>&1 | tail -2 | cut -c1-400
ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:gram_ws_kernel -s 1 -c 1 -o gpurun_out/r_gram_ws python tools/bench_configs.py --config 4 --rows 8000000 --reps 1 --gram-tier fp64 > gpurun_out/r_ncu.log 2>&1
echo "ncu rc=$?"
