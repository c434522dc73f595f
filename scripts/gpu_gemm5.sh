#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_gemm.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gemm.log
tail -12 gpurun_out/pytest_gemm.log | cut -c1-250
timeout 300 python scripts/gemm_bench.py fixed > gpurun_out/gemm_fixed_tma.log 2>&1; cut -c1-36,50-75,100-125 gpurun_out/gemm_fixed_tma.log
timeout 300 python scripts/gemm_bench.py sk > gpurun_out/gemm_bench_sk.log 2>&1; cut -c1-36,50-75,86-125 gpurun_out/gemm_bench_sk.log
