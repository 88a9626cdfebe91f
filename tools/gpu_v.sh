#!/bin/bash
# GPU call V: rehearsal of the round-end tiers on a fresh box: GPU tests, smoke, reference arm, bench (default flags)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 900 python bench.py --impl reference ) > gpurun_out/v_bench_reference.json 2> gpurun_out/v_bench_reference.err; tail -3 gpurun_out/v_bench_reference.err; cut -c1-300 gpurun_out/v_bench_reference.json
( time timeout 900 python bench.py ) > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; tail -3 gpurun_out/v_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/v_bench.json').read().strip().splitlines()[-1])
x=d.pop('extra'); print(json.dumps(d)[:1500])
print('config3', json.dumps(x['config3'])[:700])
print('config4', json.dumps(x['config4'])[:900])
print('config5', json.dumps(x['config5'])[:900])
PY
