#!/bin/bash
# weight-gradient side stream: GPU suite, then conformer / base with the stream on and off on the same box
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rf -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-300
pick() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r=d.get('roofline') or {}
print(sys.argv[1], 'trials/s', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'gemm ms', r.get('gemm_ms_per_step'), 'frac', r.get('frac'))
PY
}
B="--no-eager --no-cpu-baseline --no-inference"
timeout 600 python bench.py $B > gpurun_out/bench_conformer.log 2>&1; pick gpurun_out/bench_conformer.log
B2T_WGRAD_STREAM=0 timeout 600 python bench.py $B > gpurun_out/bench_conformer_noside.log 2>&1; pick gpurun_out/bench_conformer_noside.log
timeout 600 python bench.py --workload base $B > gpurun_out/bench_base.log 2>&1; pick gpurun_out/bench_base.log
B2T_WGRAD_STREAM=0 timeout 600 python bench.py --workload base $B > gpurun_out/bench_base_noside.log 2>&1; pick gpurun_out/bench_base_noside.log
