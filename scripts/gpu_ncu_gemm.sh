#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python scripts/gemm_bench.py epi1 > gpurun_out/gemm_epi1_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 6 -c 4 -f -o gpurun_out/gemm_epi1 \
    python scripts/gemm_bench.py epi1 > gpurun_out/gemm_epi1_ncu.log 2>&1
cat gpurun_out/gemm_epi1_plain.log; tail -3 gpurun_out/gemm_epi1_ncu.log
