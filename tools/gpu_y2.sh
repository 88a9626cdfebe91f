#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/y2_fail.txt
cat gpurun_out/y2_fail.txt | cut -c1-250
