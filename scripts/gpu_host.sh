#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_aux_gpu.py tests/test_gemm_gpu.py tests/test_models_gpu.py -m gpu -q -x -rf > gpurun_out/pytest_host.log 2>&1; tail -5 gpurun_out/pytest_host.log | cut -c1-300
python scripts/host_profile.py conformer > gpurun_out/host_profile_conformer.log 2>&1; head -48 gpurun_out/host_profile_conformer.log | cut -c1-170
python scripts/host_profile.py base 2>&1 | head -4
