#!/bin/bash
# 8-GPU conformer-large step with the round-2 data-parallel defaults (dynamic GEMM schedule, high-priority NCCL, lazy optimizer waits)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1 | sed 's/"unit.*"ms_per_step"/ms/'; grep -o '"exposed_ms_per_step": [0-9.]*, "ms_per_step_without_allreduce": [0-9.]*' $1 | head -1; }
F="--steps 12 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference"
for i in 1 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2951$i bench.py --gpus 8 $F > gpurun_out/scale8_final_$i.log 2>&1
  echo "N=8 run $i rc=$?: $(pick gpurun_out/scale8_final_$i.log)"
done
