#!/bin/bash
# round-2 check C: full GPU parity suite, smoke, the three bench workloads, conformer + base launch lists
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-2500
timeout 600 python bench.py --workload base --no-eager > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-900
timeout 600 python bench.py --workload frozen --no-eager --no-cpu-baseline > gpurun_out/bench_frozen.log 2>&1; tail -1 gpurun_out/bench_frozen.log | cut -c1-900
for w in conformer base; do
  B="python bench.py --workload $w --steps 2 --warmup 3 --no-roofline --no-cpu-baseline --no-eager --no-inference"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_$w.csv \
      $B > gpurun_out/launches_ncu_$w.log 2>&1
done
python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; tail -40 gpurun_out/kernel_bench.log
