#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in conformer; do
  B="python bench.py --workload $w --steps 2 --warmup 3 --no-roofline --no-cpu-baseline --no-eager --no-inference"
  $B > gpurun_out/plain_$w.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_$w.csv \
      $B > gpurun_out/launches_ncu_$w.log 2>&1
  python profiles/summarize_launches.py gpurun_out/launches_$w.csv | head -45
done
