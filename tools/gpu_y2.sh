#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fit.py -m gpu -x -q -k "by_column_bucket" 2>&1 | tail -3
