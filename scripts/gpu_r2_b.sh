#!/bin/bash
# round-2 check B: GPU suite + quick conformer/base bench with the arena + fused Adam
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python bench.py --no-eager --no-cpu-baseline > gpurun_out/bench_conformer_b.log 2>&1; tail -3 gpurun_out/bench_conformer_b.log | cut -c1-900
timeout 900 python bench.py --no-eager --no-cpu-baseline --optimizer torch > gpurun_out/bench_conformer_b_torchadam.log 2>&1; tail -1 gpurun_out/bench_conformer_b_torchadam.log | cut -c1-400
timeout 600 python bench.py --workload base --no-eager --no-cpu-baseline > gpurun_out/bench_base_b.log 2>&1; tail -1 gpurun_out/bench_base_b.log | cut -c1-400
