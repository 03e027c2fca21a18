#!/bin/bash
mkdir -p gpurun_out
T=${1:-final}
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/${T}_pytest.txt; tail -4 gpurun_out/${T}_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python tools/ea_quick.py 1184 2>&1 | tail -1 | tee gpurun_out/${T}_ea_quick.txt
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 300 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'], d['roofline']['frac'], d['roofline']['traffic'], d['roofline_hbm']['frac'], d['config']['other_precision_rank0'], d.get('cpu_baseline'))
PY
