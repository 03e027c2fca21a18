#!/bin/bash
mkdir -p gpurun_out
T=${1:-r3d}
timeout 900 python bench.py --workload config2 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config2.json 2> gpurun_out/${T}_bench_config2.err; tail -c 300 gpurun_out/${T}_bench_config2.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_config2.json').read().strip().splitlines()[-1])
print('config2', {k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches')}, d['e2e'], d['config']['work_items_by_kernel_rank0'])
PY
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)1, \(bool\)1, \(int\)2048" -c 1 -f -o gpurun_out/${T}_wide64_hd_2048 \
    python bench.py --workload config5 --reads 592 --steps 1 --warmup 1 > gpurun_out/${T}_ncu_hd.log 2>&1
ls -la gpurun_out/*.ncu-rep | tail -3
timeout 600 python tools/wide_quick.py contract 2368 2>&1 | tail -1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read())
print('fp64 contract 2368', 'gcups %.2f'%d['gcups'], 'ms', [round(m,1) for m in d['ms']], d['wide_items'], d['pairs'])"
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_config5.py -m gpu -x -q 2>&1 | tail -3
