#!/bin/bash
# GPU call N (2 GPUs): fence-free (tagged-word) epilogue: full suite incl. the NCCL 2-rank test, bench N=1 and N=2, step overhead
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q -x ) > gpurun_out/n_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/n_pytest.log
tail -12 gpurun_out/n_pytest.log
timeout 300 python tools/step_overhead.py > gpurun_out/n_overhead.json 2>&1; cat gpurun_out/n_overhead.json
( timeout 600 python bench.py --no-extra --no-cpu-baseline ) > gpurun_out/n_bench_n1.json 2> gpurun_out/n_bench_n1.err
cut -c1-330 gpurun_out/n_bench_n1.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
( timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-extra ) > gpurun_out/n_bench_n2.json 2> gpurun_out/n_bench_n2.err
tail -3 gpurun_out/n_bench_n2.err | cut -c1-300; cut -c1-330 gpurun_out/n_bench_n2.json
