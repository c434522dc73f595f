#!/bin/bash
# compute-sanitizer, ONE tool per gpurun call ($1 = memcheck | racecheck), on small cases of the cluster / pair kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tool=${1:-memcheck}
if [ "$tool" = memcheck ]; then
  sel='tests/test_gemm_gpu.py -k "pair_full_epilogue and 260 or pair_batched or pre_mode and 768-silu-2 or split_k_atomic"'
else
  sel='tests/test_kernels_gpu.py -k gru'
fi
timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python -m pytest $(eval echo $sel) -m gpu -q -x > gpurun_out/sanitizer_$tool.log 2>&1
echo "rc=$?" >> gpurun_out/sanitizer_$tool.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|rc=|Error|hazard" gpurun_out/sanitizer_$tool.log | head -20
