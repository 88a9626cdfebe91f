#!/bin/bash
# GPU call: dense load phases for full buckets (8 / 16 / 32 table columns): parity, sweeps
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/bucket_sweep.py --rows 200000000 --features 8,12,16,24,32 > gpurun_out/y_sweep_a.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 7,8,15,16,31 --weighted > gpurun_out/y_sweep_w.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 16,32 --loss huber > gpurun_out/y_sweep_h.jsonl
