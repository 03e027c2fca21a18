#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --strong --reads 4736 --steps 2 --warmup 3 --no-dense --no-other-precision > gpurun_out/strong_n$N.json 2> gpurun_out/strong_n$N.err; tail -c 300 gpurun_out/strong_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/strong_n$N.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','gcups','ms_per_step','scaling','n_gpus')}, d['e2e']['gcups'], d['config']['strong_scaling_shards'])
PY
