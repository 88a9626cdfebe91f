#!/bin/bash
set -x
mkdir -p gpurun_out
python tools/fuzz_passes.py 120 7 2>&1 | tail -3 | tee gpurun_out/y_fuzz.json
python tools/fuzz_passes.py 120 11 2>&1 | tail -3 | tee -a gpurun_out/y_fuzz.json
