#!/bin/bash
# GPU call K (8 GPUs): bench at N=8 with the extra block (config 3 strong-scaled, config 4, config 5), reference arm, N=4
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
( time timeout 900 $TR8 bench.py --gpus 8 --steps 20 --warmup 3 ) > gpurun_out/k_bench_n8.json 2> gpurun_out/k_bench_n8.err
echo "bench n8 rc=$?"; tail -c 800 gpurun_out/k_bench_n8.err; cut -c1-500 gpurun_out/k_bench_n8.json
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522"
( time timeout 900 $TR4 bench.py --gpus 4 --steps 20 --warmup 3 --no-extra ) > gpurun_out/k_bench_n4.json 2> gpurun_out/k_bench_n4.err
echo "bench n4 rc=$?"; cut -c1-400 gpurun_out/k_bench_n4.json
( DQ4ML_XRANK=nccl timeout 600 $TR8 bench.py --gpus 8 --steps 20 --warmup 3 --no-extra ) > gpurun_out/k_bench_n8_nccl.json 2> gpurun_out/k_bench_n8_nccl.err
cut -c1-400 gpurun_out/k_bench_n8_nccl.json
