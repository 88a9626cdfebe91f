#!/bin/bash
# GPU call: full-bucket and narrower dense load phases, final dispatch: parity, sweeps
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/bucket_sweep.py --rows 200000000 --features 8,9,10,12,14,16,17,20,24,28,32 > gpurun_out/y_sweep_final.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 7,8,12,15,16,24,31 --weighted > gpurun_out/y_sweep_final_w.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 9,12,16,24,32 --loss huber > gpurun_out/y_sweep_final_h.jsonl
