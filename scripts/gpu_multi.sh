#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
for w in base conformer; do
  timeout 240 $T bench.py --gpus 2 --workload $w > gpurun_out/bench_${w}_2gpu.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/bench_${w}_2gpu.log | cut -c1-330
done
timeout 240 $T bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench_reference_2gpu.log 2>&1; echo "ref rc=$?"; grep '^{' gpurun_out/bench_reference_2gpu.log | cut -c1-200
