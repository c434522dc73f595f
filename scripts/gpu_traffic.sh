#!/bin/bash
# DRAM bytes of every GEMM launch of one training step (ncu, dram__bytes_*), per workload -> gpurun_out/gemm_traffic_*.csv
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for w in conformer base; do
  B="python bench.py --workload $w --steps 1 --warmup 3 --no-roofline --no-cpu-baseline --no-eager --no-inference"
  $B > gpurun_out/bench_plain2_$w.log 2>&1 &&
  timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
      -k regex:gemm_tc_kernel -s 3000 -c 2400 --csv --log-file gpurun_out/gemm_traffic_$w.csv $B > gpurun_out/gemm_traffic_ncu_$w.log 2>&1
done
ls -la gpurun_out | grep traffic
