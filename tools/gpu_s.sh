#!/bin/bash
# GPU call S: the committed build: huber tests three times, then the whole suite
set -x
mkdir -p gpurun_out
for i in 1 2 3; do timeout 900 python -m pytest tests -m gpu -x -q -k "huber" 2>&1 | tail -1; done
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
python tools/bucket_sweep.py --rows 200000000 --features 9,12,17,24 --loss huber | cut -c1-260
