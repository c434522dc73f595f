#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -rf -k "gru" > gpurun_out/pytest_gru.log 2>&1; tail -4 gpurun_out/pytest_gru.log | cut -c1-250
timeout 120 python -c "
import sys; sys.path.insert(0,'scripts'); sys.argv=['x']
import kernel_bench as k; k.gru(32); k.gru(64); k.gru(96); k.gru(128)"
echo "--- B2T_GRU_MT=1"
B2T_GRU_MT=1 timeout 120 python -c "
import sys; sys.path.insert(0,'scripts'); sys.argv=['x']
import kernel_bench as k; k.gru(64)"
