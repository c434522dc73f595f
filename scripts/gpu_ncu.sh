#!/bin/bash
# One gpurun call: ncu captures (launch list of the default bench, full capture of the GEMM in isolation and in-step).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python scripts/gemm_bench.py prof > gpurun_out/gemm_prof_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -c 16 -f -o gpurun_out/gemm_iso \
    python scripts/gemm_bench.py prof > gpurun_out/gemm_iso_ncu.log 2>&1
B="python bench.py --steps 2 --warmup 3 --no-roofline --no-cpu-baseline"
$B > gpurun_out/bench_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_base.csv \
    $B > gpurun_out/launches_ncu.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:gemm_tc_kernel -s 618 -c 206 -f -o gpurun_out/gemm_step \
    $B > gpurun_out/gemm_step_ncu.log 2>&1
ls -la gpurun_out
