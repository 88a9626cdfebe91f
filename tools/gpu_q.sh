#!/bin/bash
# GPU call Q: whole GPU suite with the producer-warp fp64 Gram kernel, config 4 timings, full bench line
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
DQ4ML_GRAM_DEBUG=1 timeout 300 python tools/bench_configs.py --config 4 --reps 3 --gram-tier fp64 > gpurun_out/q_tmp.log 2>&1
grep "gram_ws pair" gpurun_out/q_tmp.log | awk '{print $3,$4,$6,$8,$9}' | sort -k1,1n -k5,5n | awk '{a[$1]=$0} END{for(k in a) print a[k]}' | sort -n
grep gram_pass_ms gpurun_out/q_tmp.log | cut -c1-420 | tee gpurun_out/q_config4_fp64.json
( timeout 600 python bench.py --no-cpu-baseline ) > gpurun_out/q_bench_n1.json 2> gpurun_out/q_bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/q_bench_n1.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['roofline']['frac'], json.dumps(d['extra']['config4'])[:900])
PY
