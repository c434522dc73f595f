#!/bin/bash
# pair GEMM + TMA-store epilogue inside the step: GPU suite, then the benches
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python bench.py --no-eager --no-cpu-baseline > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-1800
timeout 600 python bench.py --workload base --no-eager --no-cpu-baseline > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-1800
B2T_GEMM_CG=1 timeout 900 python bench.py --no-eager --no-cpu-baseline --no-inference > gpurun_out/bench_conformer_cg1.log 2>&1; tail -1 gpurun_out/bench_conformer_cg1.log | cut -c1-400
