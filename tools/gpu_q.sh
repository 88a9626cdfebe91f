#!/bin/bash
# GPU call Q: producer-warp fp64 Gram kernel: parity tests, then config 4 default split, pacing on / off
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fit.py tests/test_gpu_weighted.py tests/test_gpu_multirank.py -m gpu -x -q -k "gram or config4 or wide or weighted or multirank" 2>&1 | tail -8
for pace in 0 256 1024; do
  echo "== pace $pace"
  DQ4ML_GRAM_PACE=$pace DQ4ML_GRAM_DEBUG=1 timeout 300 python tools/bench_configs.py --config 4 --reps 3 --gram-tier fp64 > gpurun_out/q_tmp.log 2>&1
  grep "gram_ws pair" gpurun_out/q_tmp.log | awk '{print $3,$4,$6,$8,$9}' | sort -k1,1n -k5,5n | awk '{a[$1]=$0} END{for(k in a) print a[k]}' | sort -n
  grep gram_pass_ms gpurun_out/q_tmp.log | grep -o '"gram_pass_ms": [0-9.]*' | tee gpurun_out/q_pace_$pace.txt
done
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv
