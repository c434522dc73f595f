#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 240 $T scripts/dp_check.py conformer > gpurun_out/dp_check_conformer.log 2>&1; echo "dp_check conformer rc=$?"; grep "^rank" gpurun_out/dp_check_conformer.log
timeout 240 $T scripts/dp_check.py base > gpurun_out/dp_check_base.log 2>&1; echo "dp_check base rc=$?"; grep "^rank" gpurun_out/dp_check_base.log
for w in base conformer; do
  timeout 240 $T bench.py --gpus 2 --workload $w --steps 10 --warmup 3 > gpurun_out/bench_${w}_2gpu.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/bench_${w}_2gpu.log | cut -c1-330
done
