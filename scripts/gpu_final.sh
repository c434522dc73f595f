#!/bin/bash
# round-end check: GPU parity tests, smoke, default bench (both arms), conformer bench, one ncu --set full of the GEMM
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2>&1; tail -1 gpurun_out/bench_reference.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-300
timeout 600 python bench.py --workload conformer > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-300
