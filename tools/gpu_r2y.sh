#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2y}
run() { tag=$1; shift; timeout 900 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['kernel_items'], d['wide_items'], d['pairs'])" || tail -3 gpurun_out/${T}_q_${tag}.txt; }
for V in B C A; do
  cp tools/libsagpu_$V.so signalalign_b200/libsagpu.so
  run ${V}_fp64 contract 2368
done
timeout 1500 python bench.py --workload config5 --reads 3072 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config5_3072.json 2> gpurun_out/${T}_bench_config5_3072.err; tail -c 300 gpurun_out/${T}_bench_config5_3072.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_config5_3072.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'])
PY
