#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2z}
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 400 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'], d['roofline']['frac'], d['roofline']['traffic'], d['config']['other_precision_rank0'])
PY
timeout 1500 python bench.py --workload config5 --reads 3072 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config5_3072.json 2> gpurun_out/${T}_bench_config5_3072.err; tail -c 300 gpurun_out/${T}_bench_config5_3072.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_config5_3072.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'])
PY
