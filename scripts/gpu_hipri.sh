#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1 | sed 's/"unit.*"ms_per_step"/ms/'; grep -o '"exposed_ms_per_step": [0-9.]*, "ms_per_step_without_allreduce": [0-9.]*' $1 | head -1; }
F="--gpus 2 --steps 15 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference"
for c in "1 0" "0 0" "1 1" "1 0" "0 0" "1 1"; do set -- $c
  B2T_NCCL_HIPRI=$1 B2T_GEMM_CLC=$2 timeout 300 $T bench.py $F > gpurun_out/bench2_h$1c$2.log 2>&1; echo "HIPRI=$1 CLC=$2 rc=$? $(pick gpurun_out/bench2_h$1c$2.log)"
done
