#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in base conformer; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $w --steps 10 --warmup 3 > gpurun_out/bench_${w}_2gpu.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/bench_${w}_2gpu.log | cut -c1-1200
done
