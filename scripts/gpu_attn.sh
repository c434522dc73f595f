#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_kernels.log 2>&1; tail -4 gpurun_out/pytest_kernels.log | cut -c1-250
python scripts/kernel_bench.py attn 2>&1 | tee gpurun_out/attn_plain.log
