#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
pick() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'trials/s', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1))
PY
}
B="--no-eager --no-cpu-baseline --no-inference --no-roofline"
for v in 1 0 1 0; do
  B2T_GEMM_CLC=$v timeout 600 python bench.py $B > gpurun_out/bench_clc$v.log 2>&1; pick gpurun_out/bench_clc$v.log
done
for v in 1 0; do
  B2T_GEMM_CLC=$v timeout 600 python bench.py --workload base $B > gpurun_out/bench_base_clc$v.log 2>&1; pick gpurun_out/bench_base_clc$v.log
done
