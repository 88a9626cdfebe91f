#!/bin/bash
# GPU call H (2 GPUs): NCCL-bound ranks -- mailbox epilogue parity test, bench at N=2 with and without the mailboxes, reference arm under torchrun
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q > gpurun_out/h_multirank.log 2>&1
echo "multirank rc=$?"; tail -15 gpurun_out/h_multirank.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
( time timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 3 ) > gpurun_out/h_bench_n2.json 2> gpurun_out/h_bench_n2.err
echo "bench n2 rc=$?"; tail -c 1500 gpurun_out/h_bench_n2.err; cut -c1-700 gpurun_out/h_bench_n2.json
( DQ4ML_XRANK=nccl timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-extra ) > gpurun_out/h_bench_n2_nccl.json 2> gpurun_out/h_bench_n2_nccl.err
echo "bench n2 nccl rc=$?"; cut -c1-400 gpurun_out/h_bench_n2_nccl.json
( time timeout 600 $TR bench.py --impl reference --gpus 2 --steps 5 --warmup 1 ) > gpurun_out/h_bench_ref_n2.json 2> gpurun_out/h_bench_ref_n2.err
cut -c1-300 gpurun_out/h_bench_ref_n2.json; python -c "
import json
d=json.loads(open('gpurun_out/h_bench_ref_n2.json').read().strip().splitlines()[-1]); print(d['cpu_baseline']['cores'], d['value'])"
( timeout 600 python bench.py --no-extra --no-cpu-baseline ) > gpurun_out/h_bench_n1.json 2> gpurun_out/h_bench_n1.err
cut -c1-400 gpurun_out/h_bench_n1.json
