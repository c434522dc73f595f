#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python scripts/gemm_bench.py prof > gpurun_out/gemm_prof_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -c 16 -f -o gpurun_out/gemm_iso \
    python scripts/gemm_bench.py prof > gpurun_out/gemm_iso_ncu.log 2>&1
tail -3 gpurun_out/gemm_iso_ncu.log; ls -la gpurun_out
