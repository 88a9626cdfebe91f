#!/bin/bash
# GPU call Z2: ncu full capture of the 16-column loss/gradient kernel (compile-time load phase)
set -x
mkdir -p gpurun_out
python tools/pass_probe.py 16 200000000
ncu --set full --section WarpStateStats --clock-control none -k regex:stream_reduce_kernel -s 2 -c 1 -o gpurun_out/z_lsq_d16 python tools/pass_probe.py 16 200000000 > gpurun_out/z_ncu.log 2>&1
echo "ncu rc=$?"
python profiles/summarize_ncu.py gpurun_out/z_lsq_d16.ncu-rep lsq_d16_lab r2 | tail -9
cp profiles/r2_lsq_d16_lab_full.json gpurun_out/z_r2_lsq_d16_lab_full.json; cp profiles/ncu_summary.json gpurun_out/z_ncu_summary.json
rm -f gpurun_out/*.ncu-rep
