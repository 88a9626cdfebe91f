#!/bin/bash
# GPU call X: tcgen05 Gram tier: rows per fp32 accumulation span (drain interval) against time and against the fp64 tier's coefficients
set -x
mkdir -p gpurun_out
for dr in 2048 4096 8192 16384; do
  echo "== drain_rows $dr"
  DQ4ML_TF32_DRAIN_ROWS=$dr python bench.py --config 4 --no-cpu-baseline --no-extra --steps 3 --warmup 1 2>/dev/null | tail -1 > gpurun_out/x_tf32_drain_$dr.json
  python - <<PY
import json
d=json.loads(open('gpurun_out/x_tf32_drain_$dr.json').read())
f=d['fit']
print($dr, 'tf32 pass_ms', f['tf32x3']['pass_ms'], 'frac_hbm', f['tf32x3'].get('frac_of_hbm'), 'coef diff vs fp64 (scale-rel)', f['tf32x3'].get('max_coef_diff_vs_fp64_scale_rel'), 'fp64 pass', f['fp64']['pass_ms'])
PY
done
