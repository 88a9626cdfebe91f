#!/bin/bash
# GPU call U (8 GPUs): final scaling record of round 2: bench at N = 8 (with the extra block), 4, 2, 1 (no extra), 2-GPU parity tests
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
run() {  # n port extra-flags
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2"
  ( timeout 900 $TR bench.py --gpus $1 --steps 20 --warmup 3 --no-cpu-baseline $3 ) > gpurun_out/u_bench_n$1.json 2> gpurun_out/u_bench_n$1.err
  echo "bench n$1 rc=$?"; tail -1 gpurun_out/u_bench_n$1.json | cut -c1-420
}
run 8 29531 ""
run 4 29532 "--no-extra"
run 2 29533 "--no-extra"
( timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra ) > gpurun_out/u_bench_n1.json 2> gpurun_out/u_bench_n1.err
tail -1 gpurun_out/u_bench_n1.json | cut -c1-420
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q 2>&1 | tail -2
