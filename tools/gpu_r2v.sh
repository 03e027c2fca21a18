#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2v}
run() { tag=$1; shift; timeout 900 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['kernel_items'], d['wide_items'], d['pairs'])" || tail -3 gpurun_out/${T}_q_${tag}.txt; }
run long_fp64 long_runs_trim14 1184
run long_fp64_126 long_runs_trim14 1184 SA_REG_MAX_SPAN=126
run long_fp64_62 long_runs_trim14 1184 SA_REG_MAX_SPAN=62
run long_fp32_126 long_runs_trim14 1184 SA_PRECISION=fp32 SA_REG_MAX_SPAN=126
run long_fp32_62 long_runs_trim14 1184 SA_PRECISION=fp32 SA_REG_MAX_SPAN=62
run dense_fp32_126 dense 1184 SA_PRECISION=fp32 SA_REG_MAX_SPAN=126
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 600 gpurun_out/${T}_bench.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2v_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'], d['roofline']['frac'], d['config']['other_precision_rank0'], d['config']['dense_pattern_rank0'], d['config']['expectations_e2e_rank0'], d.get('cpu_baseline'))
PY
