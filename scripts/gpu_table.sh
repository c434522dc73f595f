#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
B2T_GEMM_TABLE=gpurun_out/gemm_table_conformer.txt timeout 600 python bench.py --workload conformer --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-200
B2T_GEMM_TABLE=gpurun_out/gemm_table_base.txt timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-200
