#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python scripts/kernel_bench.py attn > gpurun_out/attn_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 6 -c 2 -o gpurun_out/prof_attn python scripts/kernel_bench.py attn > gpurun_out/attn_ncu.log 2>&1
cat gpurun_out/attn_plain.log; tail -2 gpurun_out/attn_ncu.log
