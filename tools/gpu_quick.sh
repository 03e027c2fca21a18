#!/bin/bash
# quick GPU visit: parity tests + one bench line.  usage: bash tools/gpu_quick.sh <tag> [bench args]
TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.txt; tail -3 gpurun_out/${TAG}_pytest.txt
python bench.py --no-cpu --steps 2 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -3 gpurun_out/${TAG}_bench.err
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench.json')); print('gcups', d['gcups'], 'ms', d['ms_per_step'], 'e2e gcups', d['e2e']['gcups'], 'frac', d['roofline']['frac'])"
