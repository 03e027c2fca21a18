#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2g}
nproc > gpurun_out/${T}_nproc.txt
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${T}_bench.err
cut -c1-3000 gpurun_out/${T}_bench.json
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; cut -c1-700 gpurun_out/${T}_bench_reference.json
