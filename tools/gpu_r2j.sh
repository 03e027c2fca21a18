#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2j}
python bench.py --workload config5 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config5.json 2> gpurun_out/${T}_bench_config5.err; echo "c5 exit $?"; tail -2 gpurun_out/${T}_bench_config5.err; cut -c1-1800 gpurun_out/${T}_bench_config5.json
python bench.py --workload config4 --steps 2 --warmup 1 --reads 2368 --no-cpu > gpurun_out/${T}_bench_config4.json 2> gpurun_out/${T}_bench_config4.err; echo "c4 exit $?"; tail -2 gpurun_out/${T}_bench_config4.err; cut -c1-1800 gpurun_out/${T}_bench_config4.json
