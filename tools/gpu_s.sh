#!/bin/bash
# GPU call S: lane-group kernel shapes for 17..24 columns and the per-pass sweep over all buckets
set -x
mkdir -p gpurun_out
for sh in 0 1 2; do
  DQ4ML_PAIR24_SHAPE=$sh timeout 600 python -m pytest tests/test_gpu_fit.py -m gpu -x -q -k "summarizer_and_lsq" 2>&1 | tail -1
  DQ4ML_PAIR24_SHAPE=$sh python tools/bucket_sweep.py --rows 200000000 --features 17,18,20,22,24 | sed "s/^{/{\"shape24\": $sh, /" | tee -a gpurun_out/s_shapes24.jsonl | cut -c1-330
done
python tools/bucket_sweep.py --rows 200000000 --features 2,4,8,12,16,25,28,32,40,70 | tee gpurun_out/s_bucket_sweep.jsonl | cut -c1-330
python tools/bucket_sweep.py --rows 200000000 --features 8,16,17,24,32 --weighted | tee gpurun_out/s_bucket_sweep_weighted.jsonl | cut -c1-330
