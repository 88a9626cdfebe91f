#!/bin/bash
# GPU call J: TMA staging of the tf32 Gram kernel, L2 promotion variants vs cp.async, + ncu of the TMA variant
set -x
mkdir -p gpurun_out
for l2 in 128 0 256; do
  echo "{\"tma_l2_promotion\": $l2}" >> gpurun_out/j_config4.jsonl
  DQ4ML_TF32_TMA_L2=$l2 timeout 300 python tools/bench_configs.py --config 4 --gram-tier tf32x3 --reps 3 2>&1 | cut -c1-300 | grep gram_pass_ms >> gpurun_out/j_config4.jsonl
done
echo '{"tma": "off (cp.async)"}' >> gpurun_out/j_config4.jsonl
DQ4ML_TF32_TMA=0 timeout 300 python tools/bench_configs.py --config 4 --gram-tier tf32x3 --reps 3 2>&1 | cut -c1-300 | grep gram_pass_ms >> gpurun_out/j_config4.jsonl
cut -c1-200 gpurun_out/j_config4.jsonl
timeout 300 python tools/bench_configs.py --config 4 --gram-tier tf32x3 --rows 8000000 --reps 1 > gpurun_out/j_plain.log 2>&1 && \
timeout 900 ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:gram_tf32_kernel -s 1 -c 1 -o gpurun_out/j_tf32_tma python tools/bench_configs.py --config 4 --gram-tier tf32x3 --rows 8000000 --reps 1 > gpurun_out/j_ncu.log 2>&1
echo "ncu rc=$?"
