#!/bin/bash
# GPU call S: the extended parity tests (one per kernel family of the streaming passes), then the whole suite
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "summarizer_and_lsq or by_column_bucket or huber_fit_vs_oracle or weighted_lbfgs_fit_vs_oracle" 2>&1 | tail -15
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
cp gpurun_out/parity_errors.jsonl gpurun_out/s_parity_errors.jsonl
