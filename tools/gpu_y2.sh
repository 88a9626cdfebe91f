#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fit.py -m gpu -x -q -k "tile_boundaries or wls_moments or lab_app or config2" 2>&1 | tail -15 | cut -c1-250
