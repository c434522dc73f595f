#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/gemm_bench.py epi1 > gpurun_out/epi1_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 4 -c 1 -o gpurun_out/prof_gemm_epi_dgrad python scripts/gemm_bench.py epi1 > gpurun_out/epi1_ncu.log 2>&1
cat gpurun_out/epi1_plain.log; tail -2 gpurun_out/epi1_ncu.log
