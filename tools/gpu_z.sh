#!/bin/bash
# GPU call Z: final evidence: ncu full capture of the 16-column loss/gradient kernel (compile-time load phase), then the round-end tiers
set -x
mkdir -p gpurun_out
python tools/pass_probe.py 16 200000000
ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:stream_reduce_kernel -s 2 -c 1 -o gpurun_out/z_lsq_d16 python tools/pass_probe.py 16 200000000 > gpurun_out/z_ncu.log 2>&1
echo "ncu rc=$?"
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
( timeout 900 python bench.py --no-cpu-baseline ) > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; tail -2 gpurun_out/z_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/z_bench.json').read().strip().splitlines()[-1])
x=d.pop('extra'); print(d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['clocks']['reasons'])
print('config3', x['config3']['fit_seconds'], x['config3']['summarizer_pass'], x['config3']['lsq_pass'])
print('config4', x['config4']['fp64']['pass_ms'], x['config4']['tf32x3']['pass_ms'])
print('config5', [(e['rows_per_gpu'], round(e['frac'],3)) for e in x['config5']])
PY
