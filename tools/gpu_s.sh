#!/bin/bash
# GPU call S: huber / evaluation at 17..32 columns: groups of four lanes (DQ4ML_LG_QUAD=1, default) against the generic 32-column bucket
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for qd in 1 0; do
  DQ4ML_LG_QUAD=$qd python tools/bucket_sweep.py --rows 200000000 --features 17,24,32 --loss huber | sed "s/^{/{\"quad\": $qd, /" >> gpurun_out/s_quad_huber.jsonl
  DQ4ML_LG_QUAD=$qd python tools/bucket_sweep.py --rows 200000000 --features 8,12,16,17,24,32 | sed "s/^{/{\"quad\": $qd, /" >> gpurun_out/s_quad_eval.jsonl
done
