#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2t}
run() { tag=$1; shift; timeout 900 python tools/wide_quick.py "$@" > gpurun_out/${T}_q_${tag}.txt 2>&1; tail -1 gpurun_out/${T}_q_${tag}.txt | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('$tag', d['env'], 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['wide_items'], d['pairs'])" || tail -3 gpurun_out/${T}_q_${tag}.txt; }
timeout 1500 python -m pytest tests/test_gpu_wide.py tests/test_gpu_config5.py tests/test_zz_no_alloc.py tests/test_cli_dropin.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_pytest.txt; tail -5 gpurun_out/${T}_pytest.txt
run fp64 contract 2368
run fp64_foldg contract 2368 SA_WIDE_FOLDG_FROM=1024
run fp64_4736 contract 4736
