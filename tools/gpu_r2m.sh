#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2m}
run() { tag=$1; shift; timeout 900 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['wide_items'], d['pairs'])" || tail -3 gpurun_out/${T}_q_${tag}.txt; }
for V in A B; do
  cp tools/libsagpu_$V.so signalalign_b200/libsagpu.so
  timeout 1500 python -m pytest tests/test_gpu_wide32.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${T}_pytest_$V.txt; tail -3 gpurun_out/${T}_pytest_$V.txt
  run ${V}_fp32 contract 2368 SA_PRECISION=fp32
  run ${V}_fp32_4736 contract 4736 SA_PRECISION=fp32
done
run fp64 contract 2368
