#!/bin/bash
cd "$(dirname "$0")/.."
for v in 0 1 0 1; do
  echo "--- B2T_GEMM_CLC=$v"
  B2T_GEMM_CLC=$v timeout 200 python scripts/gemm_bench.py sk 2>&1 | grep "sk=-1" | head -8 | cut -c1-36,50-75,105-135
done
