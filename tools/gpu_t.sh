#!/bin/bash
# GPU call T: ncu of the ring lane-group loss/gradient kernel at 24 columns (G = 4) and 16 columns (G = 2)
set -x
mkdir -p gpurun_out
python tools/pass_probe.py 24 100000000
python tools/pass_probe.py 16 100000000
ncu --set full --section WarpStateStats --section SourceCounters --clock-control none --import-source on -k regex:lane_group -s 2 -c 1 -o gpurun_out/t_lg_d24 python tools/pass_probe.py 24 100000000 > gpurun_out/t_ncu.log 2>&1
echo "ncu rc=$?"
