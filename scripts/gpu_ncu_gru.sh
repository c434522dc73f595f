#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/kernel_bench.py gru > gpurun_out/gru_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gru_fwd -s 3 -c 1 -o gpurun_out/prof_gru python scripts/kernel_bench.py gru > gpurun_out/gru_ncu.log 2>&1
cat gpurun_out/gru_plain.log; tail -2 gpurun_out/gru_ncu.log
