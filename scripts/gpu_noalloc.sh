#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -rf > gpurun_out/pytest_noalloc_alone.log 2>&1; tail -3 gpurun_out/pytest_noalloc_alone.log | cut -c1-250
timeout 1500 python -m pytest tests -m gpu -q -x -rf > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log | cut -c1-250
