#!/bin/bash
# GPU call W: soak of the fp64 Gram kernel (bit-identical repeats at 2 x 10^7 ragged rows), GPU suite twice more
set -x
mkdir -p gpurun_out
python tools/gram_soak.py 20000019 12 | tee gpurun_out/w_gram_soak.json
for i in 1 2; do timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -1; done
python bench.py --config 5 --no-cpu-baseline --no-extra --steps 3 --warmup 1 --rows 100000000 2>/dev/null | tail -1 | cut -c1-900
