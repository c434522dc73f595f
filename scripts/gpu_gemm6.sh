#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_gemm.log 2>&1; tail -3 gpurun_out/pytest_gemm.log | cut -c1-250
timeout 300 python scripts/gemm_bench.py epi2 > gpurun_out/gemm_epi2_tma.log 2>&1
cut -c1-75 gpurun_out/gemm_epi2_tma.log
