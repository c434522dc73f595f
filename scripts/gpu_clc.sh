#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_clc.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_clc.log | cut -c1-300
grep -m3 "b2t_gemm:" gpurun_out/pytest_clc.log
for v in 1 0; do
  echo "--- B2T_GEMM_CLC=$v"
  B2T_GEMM_CLC=$v timeout 200 python scripts/gemm_bench.py cg 2>&1 | grep "terms=3" | grep "cg=   2" | grep "sk=-1" | cut -c1-36,50-75,100-125
done
