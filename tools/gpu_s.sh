#!/bin/bash
# GPU call S: groups of four lanes, tall shape up to 24 columns for evaluation / huber / gradient groups
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "by_column_bucket or huber or wide or d40 or d70 or training_summary" 2>&1 | tail -2
python tools/bucket_sweep.py --rows 200000000 --features 17,20,24 | sed "s/^{/{\"quad\": 2, /" >> gpurun_out/s_quad_eval.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 17,24 --loss huber | sed "s/^{/{\"quad\": 2, /" >> gpurun_out/s_quad_huber.jsonl
