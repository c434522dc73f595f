#!/bin/bash
# round-2 evidence set: GPU suite, smoke, the three workloads (conformer with the eager + CPU baselines), kernel timings,
# launch lists, the GEMM's ncu --set full capture, the reference arm
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_conformer.log 2>&1; tail -1 gpurun_out/bench_conformer.log | cut -c1-300
timeout 600 python bench.py --workload base --no-eager > gpurun_out/bench_base.log 2>&1; tail -1 gpurun_out/bench_base.log | cut -c1-300
timeout 600 python bench.py --workload frozen --no-eager --no-cpu-baseline > gpurun_out/bench_frozen.log 2>&1; tail -1 gpurun_out/bench_frozen.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2>&1; tail -1 gpurun_out/bench_reference.log | cut -c1-300
python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; cat gpurun_out/kernel_bench.log
python scripts/gemm_bench.py cg > gpurun_out/gemm_bench_cg.log 2>&1
python scripts/gemm_bench.py epi2 > gpurun_out/gemm_epi2.log 2>&1
for w in conformer base; do
  B="python bench.py --workload $w --steps 2 --warmup 3 --no-roofline --no-cpu-baseline --no-eager --no-inference"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_$w.csv $B > gpurun_out/launches_ncu_$w.log 2>&1
done
python scripts/gemm_bench.py prof2 > gpurun_out/prof2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 3 -c 5 -o gpurun_out/prof_gemm_pair_final python scripts/gemm_bench.py prof2 > gpurun_out/prof2_ncu.log 2>&1
ls -la gpurun_out | tail -30
