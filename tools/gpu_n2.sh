#!/bin/bash
mkdir -p gpurun_out
T=${1:-n2}
nproc
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; tail -c 300 gpurun_out/${T}_bench_n2.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_n2.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches','n_gpus')}, d['e2e'], d['config']['expectations_e2e_rank0'], d['config']['host_threads_per_rank'])
PY
timeout 900 python -m pytest tests/test_gpu_comm.py -m gpu -x -q 2>&1 | tail -3
