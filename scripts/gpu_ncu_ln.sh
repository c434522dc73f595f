#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python scripts/kernel_bench.py ln > gpurun_out/ln_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ln_ -s 20 -c 6 -f -o gpurun_out/ln_prof \
    python scripts/kernel_bench.py ln > gpurun_out/ln_ncu.log 2>&1
cat gpurun_out/ln_plain.log; tail -2 gpurun_out/ln_ncu.log
