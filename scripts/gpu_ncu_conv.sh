#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python scripts/kernel_bench.py conv > gpurun_out/conv_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:glu_dwconv_bwd -s 5 -c 2 -f -o gpurun_out/conv_prof \
    python scripts/kernel_bench.py conv > gpurun_out/conv_ncu.log 2>&1
cat gpurun_out/conv_plain.log; tail -2 gpurun_out/conv_ncu.log
