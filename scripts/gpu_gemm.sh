#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gemm_gpu.py -x -q > gpurun_out/pytest_gemm.log 2>&1; tail -5 gpurun_out/pytest_gemm.log
timeout 600 python scripts/gemm_bench.py epi > gpurun_out/gemm_epi.log 2>&1; cat gpurun_out/gemm_epi.log
