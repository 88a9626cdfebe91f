#!/bin/bash
# GPU call Z3: the bench line with the extra block, shipped code
set -x
mkdir -p gpurun_out
( timeout 900 python bench.py --no-cpu-baseline ) > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; tail -2 gpurun_out/z_bench.err
cut -c1-200 gpurun_out/z_bench.json
