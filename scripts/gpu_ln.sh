#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_conformer_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_ln.log 2>&1; tail -4 gpurun_out/pytest_ln.log | cut -c1-250
python scripts/kernel_bench.py ln; python -c "
import sys; sys.path.insert(0,'scripts'); sys.argv=['x']
import kernel_bench as k; k.ln(7552, 768)"
echo "--- B2T_LN_BWD_FUSED=0"; B2T_LN_BWD_FUSED=0 python scripts/kernel_bench.py ln
python scripts/kernel_bench.py pre
