#!/bin/bash
cd "$(dirname "$0")/.."
for d in 0 1 2 4 3 5 6; do echo "== B2T_GEMM_DEBUG=$d (1 no epilogue, 2 no MMA, 4 no TMA)"; B2T_GEMM_DEBUG=$d timeout 120 python scripts/gemm_bench.py prof x; done
