#!/bin/bash
mkdir -p gpurun_out
T=${1:-r3a}
for R in 768 3072; do
timeout 1500 python bench.py --workload config5 --reads $R --steps 2 --warmup 1 > gpurun_out/${T}_bench_config5_$R.json 2> gpurun_out/${T}_bench_config5_$R.err; tail -c 300 gpurun_out/${T}_bench_config5_$R.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_config5_$R.json').read().strip().splitlines()[-1])
print($R, {k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'])
PY
done
