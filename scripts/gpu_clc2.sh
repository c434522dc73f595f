#!/bin/bash
cd "$(dirname "$0")/.."
for c in "1 1" "1 0" "0 1" "0 0"; do set -- $c
  echo "--- CLC=$1 PDL=$2"
  B2T_GEMM_CLC=$1 B2T_GEMM_PDL=$2 timeout 200 python scripts/gemm_bench.py fixed 2>&1 | grep "cg=   2" | grep -v "hi/lo" | cut -c1-36,50-75,100-125
done
