#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
T=n$N
nproc; free -g | head -2
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 400 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','dtype','gpu_launches','n_gpus')}, d['e2e'], d['config']['expectations_e2e_rank0'], d['config']['host_threads_per_rank'])
PY
