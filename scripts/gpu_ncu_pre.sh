#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "smooth or augment" 2>&1 | tail -1
python scripts/kernel_bench.py pre > gpurun_out/pre_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:smooth_fwd -s 3 -c 1 -o gpurun_out/prof_smooth python scripts/kernel_bench.py pre > gpurun_out/pre_ncu.log 2>&1
cat gpurun_out/pre_plain.log
