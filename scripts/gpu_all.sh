#!/bin/bash
cd "$(dirname "$0")/.."
bash scripts/gpu_table.sh
timeout 300 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; grep -i "layernorm" gpurun_out/kernel_bench.log
bash scripts/gpu_traffic.sh
