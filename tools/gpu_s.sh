#!/bin/bash
# GPU call S: warp-cooperative lane-group kernels (DQ4ML_LG_WC=1) for 17..32 columns: parity subset, then sweeps
set -x
mkdir -p gpurun_out
DQ4ML_LG_WC=1 timeout 1500 python -m pytest tests -m gpu -x -q -k "summarizer_and_lsq or wide or weighted or huber or d40 or d70 or training_summary or multirank" 2>&1 | tail -3
DQ4ML_LG_WC=1 python tools/bucket_sweep.py --rows 200000000 --features 17,20,24,25,28,32 | sed "s/^{/{\"wc\": 1, /" >> gpurun_out/s_wc_sweep.jsonl
DQ4ML_LG_WC=1 python tools/bucket_sweep.py --rows 200000000 --features 17,24,31 --weighted | sed "s/^{/{\"wc\": 1, /" >> gpurun_out/s_wc_sweep_weighted.jsonl
DQ4ML_LG_WC=1 python tools/bucket_sweep.py --rows 200000000 --features 17,24,32 --loss huber | sed "s/^{/{\"wc\": 1, /" >> gpurun_out/s_wc_sweep_huber.jsonl
DQ4ML_LG_WC=0 python tools/bucket_sweep.py --rows 200000000 --features 17,24,32 | sed "s/^{/{\"wc\": 0, /" >> gpurun_out/s_wc_sweep.jsonl
