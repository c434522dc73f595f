#!/bin/bash
# 8-GPU conformer step: SM partition between the persistent GEMM grid and NCCL
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "host cores: $(nproc)"
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1 | sed 's/"unit.*"ms_per_step"/ms/'; grep -o '"exposed_ms_per_step": [0-9.]*, "ms_per_step_without_allreduce": [0-9.]*' $1 | head -1; }
F="--steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 $F > gpurun_out/scale8_$name.log 2>&1
  echo "$name rc=$?: $(pick gpurun_out/scale8_$name.log)"
}
run sms132_cta16 B2T_GEMM_SMS=132 NCCL_MAX_CTAS=16
run sms140_cta8 B2T_GEMM_SMS=140 NCCL_MAX_CTAS=8
run sms124_cta24 B2T_GEMM_SMS=124 NCCL_MAX_CTAS=24
run noside_sms132_cta16 B2T_WGRAD_STREAM=0 B2T_GEMM_SMS=132 NCCL_MAX_CTAS=16
run plain B2T_X=0
