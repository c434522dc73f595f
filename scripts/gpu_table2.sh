#!/bin/bash
# per-shape GEMM tables of the conformer step for three kernel configurations
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --no-eager --no-cpu-baseline --no-inference --steps 10"
B2T_GEMM_CG=1 B2T_GEMM_NO_TMA_EPILOGUE=1 B2T_GEMM_TABLE=gpurun_out/table_cg1_old.txt $B > gpurun_out/bench_t1.log 2>&1; tail -1 gpurun_out/bench_t1.log | cut -c1-120
B2T_GEMM_NO_TMA_EPILOGUE=1 B2T_GEMM_TABLE=gpurun_out/table_cg2_old.txt $B > gpurun_out/bench_t2.log 2>&1; tail -1 gpurun_out/bench_t2.log | cut -c1-120
B2T_GEMM_TABLE=gpurun_out/table_cg2_tma.txt $B > gpurun_out/bench_t3.log 2>&1; tail -1 gpurun_out/bench_t3.log | cut -c1-120
B2T_GEMM_CG=1 B2T_GEMM_TABLE=gpurun_out/table_cg1_tma.txt $B > gpurun_out/bench_t4.log 2>&1; tail -1 gpurun_out/bench_t4.log | cut -c1-120
