#!/bin/bash
# GPU call Y: the software-pipelined single-column kernel: whole GPU suite, then the bench A/B against the grid-strided kernel
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for pipe in 1 0 1 0; do
  DQ4ML_STREAM_PIPE=$pipe python bench.py --no-cpu-baseline --no-extra 2>/dev/null | tail -1 > gpurun_out/y_bench_pipe$pipe.json
  python - <<PY
import json
d=json.loads(open('gpurun_out/y_bench_pipe$pipe.json').read())
print('pipe', $pipe, 'ms/step', round(d['ms_per_step'],5), 'kernel ms', round(d['roofline']['avg_launch_ms'],5), 'frac', round(d['roofline']['frac'],4), 'e2e', f"{d['e2e']['value']:.4e}", d['clocks']['reasons'], d['fit']['coefficients'] if 'coefficients' in d['fit'] else '')
PY
done
