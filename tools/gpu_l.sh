#!/bin/bash
# GPU call L: full GPU suite on the current build (wide huber, weighted wide l-bfgs, TMA equality), smoke
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q ) > gpurun_out/l_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/l_pytest.log
tail -25 gpurun_out/l_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
