#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -rf -k "gru" > gpurun_out/pytest_gru.log 2>&1; tail -4 gpurun_out/pytest_gru.log | cut -c1-250
python scripts/kernel_bench.py gru 2>&1 | tee gpurun_out/gru_plain.log
python -c "
import sys; sys.path.insert(0,'scripts'); sys.argv=['x']
import kernel_bench as k; k.gru(64); k.gru(128)"
