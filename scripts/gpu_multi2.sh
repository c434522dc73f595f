#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 240 $T scripts/dp_check.py conformer > gpurun_out/dp_check_conformer.log 2>&1; echo "dp_check conformer rc=$?"; grep "^rank" gpurun_out/dp_check_conformer.log
timeout 200 $T bench.py --gpus 2 --workload conformer --steps 10 --warmup 3 --no-roofline > gpurun_out/bench_conformer_2gpu_defer.log 2>&1
echo "defer rc=$? $(grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": 2[^}]*ms_per_step": [0-9.]*' gpurun_out/bench_conformer_2gpu_defer.log)"
timeout 200 $T bench.py --gpus 2 --workload base --steps 10 --warmup 3 --no-roofline > gpurun_out/bench_base_2gpu_defer.log 2>&1
echo "defer base rc=$? $(grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": 2[^}]*ms_per_step": [0-9.]*' gpurun_out/bench_base_2gpu_defer.log)"
