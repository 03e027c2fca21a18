#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2k}
timeout 1500 python -m pytest tests/test_gpu_wide32.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/${T}_pytest.txt; tail -12 gpurun_out/${T}_pytest.txt
run() { tag=$1; shift; timeout 900 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['wide_items'], d['pairs'])" || tail -3 gpurun_out/${T}_q_${tag}.txt; }
run fp64 contract 2368
run fp32 contract 2368 SA_PRECISION=fp32
run fp32_gr1024 contract 2368 SA_PRECISION=fp32 SA_WIDE32_GR_FROM=1024
run fp32_gr4096 contract 2368 SA_PRECISION=fp32 SA_WIDE32_GR_FROM=4096
run fp32_nt contract 2368 SA_PRECISION=fp32 SA_WIDE_NT=128,256,256,256,512
