#!/bin/bash
# One gpurun call: GPU parity tests, smoke, benches of both workloads, launch list, one full ncu capture of the GEMM.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log
timeout 600 python bench.py --workload conformer --steps 20 --warmup 5 > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log
timeout 300 python scripts/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; tail -20 gpurun_out/gemm_bench.log
timeout 300 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; tail -20 gpurun_out/kernel_bench.log
