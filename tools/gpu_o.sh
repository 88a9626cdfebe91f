#!/bin/bash
# GPU call O: why is the bench kernel ~7 us slower inside bench.py than in tools/step_overhead.py?
set -x
mkdir -p gpurun_out
run() { tag=$1; shift; ( "$@" ) > gpurun_out/o_$tag.json 2>/dev/null; python - <<PY
import json
d=json.loads(open('gpurun_out/o_$tag.json').read().strip().splitlines()[-1])
print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'kernel ms %.4f'%d['roofline']['avg_launch_ms'], d['clocks'])
PY
}
run sampler python bench.py --no-extra --no-cpu-baseline
run nosampler python bench.py --no-extra --no-cpu-baseline --no-clock-sampler
run nosampler_nopreroll python bench.py --no-extra --no-cpu-baseline --no-clock-sampler --preroll-seconds 0
DQ4ML_EPILOGUE=off run epoff_sampler python bench.py --no-extra --no-cpu-baseline
DQ4ML_EPILOGUE=off run epoff_nosampler python bench.py --no-extra --no-cpu-baseline --no-clock-sampler
python tools/step_overhead.py
