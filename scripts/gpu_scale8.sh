#!/bin/bash
# 1/2/4/8-GPU weak scaling of the conformer-large step (BASELINE configs[3]/[4]) on one box, plus the 2-rank gradient check
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
pick() { grep -o '"value": [0-9.]*, "unit": "trials/s", "n_gpus": [0-9]*[^}]*ms_per_step": [0-9.]*' $1 | head -1; grep -o '"comm": {[^}]*}' $1 | head -1; }
F="--steps 10 --warmup 3 --no-roofline --no-eager --no-cpu-baseline --no-inference"
timeout 200 python bench.py $F > gpurun_out/scale_conformer_1.log 2>&1; echo "N=1: $(pick gpurun_out/scale_conformer_1.log)"
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n $F > gpurun_out/scale_conformer_$n.log 2>&1
  echo "N=$n rc=$?: $(pick gpurun_out/scale_conformer_$n.log)"
done
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 scripts/dp_check.py conformer > gpurun_out/dp_check_conformer.log 2>&1; echo "dp_check rc=$?"; grep "^rank" gpurun_out/dp_check_conformer.log | head -2
