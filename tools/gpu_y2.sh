#!/bin/bash
# GPU call: lane-group kernels with uniform pointer loads + select
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "summarizer_and_lsq or by_column_bucket or weighted or huber" 2>&1 | tail -1
python tools/bucket_sweep.py --rows 200000000 --features 9,10,12,14 > gpurun_out/y_sweep_uni.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 8,12,16 --weighted > gpurun_out/y_sweep_uni_w.jsonl
