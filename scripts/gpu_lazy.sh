#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1 | sed 's/"unit.*"ms_per_step"/ms/'; grep -o '"exposed_ms_per_step": [0-9.]*, "ms_per_step_without_allreduce": [0-9.]*' $1 | head -1; }
F="--gpus 2 --steps 15 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference"
for v in 0 1 0 1 0 1; do
  B2T_DP_LAZY=$v timeout 300 $T bench.py $F > gpurun_out/bench2_lazy$v.log 2>&1; echo "LAZY=$v rc=$? $(pick gpurun_out/bench2_lazy$v.log)"
done
