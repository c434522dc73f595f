#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/gemm_bench.py prof2 > gpurun_out/prof2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 3 -c 5 -o gpurun_out/prof_gemm_pair python scripts/gemm_bench.py prof2 > gpurun_out/prof2_ncu.log 2>&1
tail -3 gpurun_out/prof2_plain.log; tail -3 gpurun_out/prof2_ncu.log; ls -la gpurun_out/*.ncu-rep
