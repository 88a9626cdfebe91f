#!/bin/bash
# GPU call M: memcheck of the wide huber path (one sanitizer tool in this call)
set -x
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest "tests/test_gpu_huber.py::test_huber_fit_vs_oracle" -x -q -k "70 and False" > gpurun_out/m_memcheck.log 2>&1
grep -n "Invalid\|at 0x\|by thread\|Address\|========= \+at\|in .*kernel\|ERROR SUMMARY" gpurun_out/m_memcheck.log | head -30
tail -5 gpurun_out/m_memcheck.log
