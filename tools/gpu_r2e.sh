#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2e}
timeout 1500 python -m pytest tests/test_gpu_wide.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/${T}_pytest.txt; tail -3 gpurun_out/${T}_pytest.txt
run() { tag=$1; shift; timeout 600 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['wide_items'])"; }
run base contract 1184
run gr4096 contract 1184 SA_WIDE_GR_FROM=4096 SA_WIDE_NT=128,256,512,1024,1024
run dense_wide dense 2368 SA_WIDE_ONLY=1
run base2368 contract 2368
