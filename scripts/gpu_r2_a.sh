#!/bin/bash
# round-2 check A: full GPU parity suite (incl. benchmarked configs), smoke, the three bench workloads
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-1500
timeout 600 python bench.py --workload base --no-eager > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-600
timeout 600 python bench.py --workload frozen --no-eager --no-cpu-baseline > gpurun_out/bench_frozen.log 2>&1; tail -1 gpurun_out/bench_frozen.log | cut -c1-600
