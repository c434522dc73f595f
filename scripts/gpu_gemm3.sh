#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python scripts/gemm_bench.py sk > gpurun_out/gemm_bench_sk.log 2>&1; cat gpurun_out/gemm_bench_sk.log
for d in 8; do
echo "--- B2T_GEMM_DEBUG=$d (8: no exchange at all, 16: no wait, 32: no partial data traffic; wrong results)"
B2T_GEMM_DEBUG=$d timeout 300 python scripts/gemm_bench.py sk > gpurun_out/gemm_bench_sk_dbg$d.log 2>&1; grep "sk= 1" gpurun_out/gemm_bench_sk_dbg$d.log
done
timeout 600 python scripts/gemm_bench.py cg > gpurun_out/gemm_bench_cg.log 2>&1; grep "terms=3" gpurun_out/gemm_bench_cg.log
