#!/bin/bash
# GPU call S: final lane-group dispatch: whole GPU suite, then the bucket sweeps that go to profiles/
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/bucket_sweep.py --rows 200000000 --features 2,4,8,9,10,12,14,16,17,20,24,25,28,32 > gpurun_out/s_bucket_sweep.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 4,8,10,12,15,16,24,31 --weighted > gpurun_out/s_bucket_sweep_weighted.jsonl
python tools/bucket_sweep.py --rows 200000000 --features 8,9,12,16,24,32 --loss huber > gpurun_out/s_bucket_sweep_huber.jsonl
DQ4ML_LANE_GROUPS=0 python tools/bucket_sweep.py --rows 200000000 --features 9,12 --loss huber > gpurun_out/s_bucket_sweep_huber_generic.jsonl
DQ4ML_LANE_GROUPS=0 python tools/bucket_sweep.py --rows 200000000 --features 16 --weighted > gpurun_out/s_bucket_sweep_weighted16_generic.jsonl
