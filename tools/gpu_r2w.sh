#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2w}
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_pytest.txt; tail -6 gpurun_out/${T}_pytest.txt
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 600 gpurun_out/${T}_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'], d['roofline']['frac'], d['config']['other_precision_rank0'], d['config']['dense_pattern_rank0'], d['config']['expectations_e2e_rank0'], d.get('cpu_baseline'))
PY
