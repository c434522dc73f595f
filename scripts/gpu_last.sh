#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 300 python bench.py --no-eager --no-cpu-baseline --no-inference --steps 10 2>&1 | tail -1 | cut -c1-330
