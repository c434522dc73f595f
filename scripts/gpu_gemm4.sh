#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for d in 0 1 3 7; do
echo "--- B2T_GEMM_DEBUG=$d (1: no epilogue body, 2: no MMA issue, 4: no TMA loads; forces single-CTA tiles)"
B2T_GEMM_DEBUG=$d timeout 300 python scripts/gemm_bench.py fixed 2>&1 | tee gpurun_out/gemm_fixed_dbg$d.log | cut -c1-60,100-140
done
