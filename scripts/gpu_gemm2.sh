#!/bin/bash
# pair-mode GEMM: parity tests, then timings of the step's shapes with single-CTA and pair tiles
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_gemm.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gemm.log
tail -15 gpurun_out/pytest_gemm.log | cut -c1-250
timeout 600 python scripts/gemm_bench.py cg > gpurun_out/gemm_bench_cg.log 2>&1; tail -60 gpurun_out/gemm_bench_cg.log
