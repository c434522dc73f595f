#!/bin/bash
# 2-GPU data parallel: gradient check, conformer (configs[4]) and base, early path on / off, 1-GPU line on the same box
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1; }
timeout 300 $T scripts/dp_check.py conformer > gpurun_out/dp_check_conformer.log 2>&1; echo "dp_check conformer rc=$?"; grep "^rank" gpurun_out/dp_check_conformer.log | head -4
timeout 300 python bench.py --steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference > gpurun_out/bench_conformer_1gpu_m.log 2>&1; echo "1 gpu: $(pick gpurun_out/bench_conformer_1gpu_m.log)"
timeout 300 $T bench.py --gpus 2 --steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference > gpurun_out/bench_conformer_2gpu.log 2>&1
echo "2 gpu early rc=$? $(pick gpurun_out/bench_conformer_2gpu.log)"
B2T_DP_EARLY=0 timeout 300 $T bench.py --gpus 2 --steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference > gpurun_out/bench_conformer_2gpu_late.log 2>&1
echo "2 gpu late rc=$? $(pick gpurun_out/bench_conformer_2gpu_late.log)"
timeout 300 $T bench.py --gpus 2 --workload base --steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference > gpurun_out/bench_base_2gpu.log 2>&1
echo "2 gpu base rc=$? $(pick gpurun_out/bench_base_2gpu.log)"
tail -3 gpurun_out/bench_conformer_2gpu.log | cut -c1-1500
