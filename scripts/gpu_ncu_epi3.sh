#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/gemm_bench.py epi3 > gpurun_out/epi3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 6 -c 4 -o gpurun_out/prof_gemm_epi3 python scripts/gemm_bench.py epi3 > gpurun_out/epi3_ncu.log 2>&1
cat gpurun_out/epi3_plain.log
