#!/bin/bash
# ncu launch lists (gpu__time_duration per launch) of both workloads' training steps
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for w in base conformer; do
  B="python bench.py --workload $w --steps 2 --warmup 3 --no-roofline --no-cpu-baseline"
  $B > gpurun_out/bench_plain_$w.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_$w.csv \
      $B > gpurun_out/launches_ncu_$w.log 2>&1
  tail -1 gpurun_out/bench_plain_$w.log | cut -c1-160
done
ls -la gpurun_out | grep launches
