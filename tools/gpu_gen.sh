#!/bin/bash
mkdir -p gpurun_out
T=${1:-r3n}
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_shim.py tests/test_cli_dropin.py tests/test_gpu_wide.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py --workload config2 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config2.json 2> gpurun_out/${T}_bench_config2.err; tail -c 300 gpurun_out/${T}_bench_config2.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_config2.json').read().strip().splitlines()[-1])
print('config2', {k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e']['gcups'], d['config']['aligned_pairs_rank0'], d['config']['pairs_column_sums_rank0'])
PY
timeout 600 python tools/config2.py 592 2>&1 | tail -2
