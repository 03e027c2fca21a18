#!/bin/bash
# one full ncu capture of the CTA kernel.  usage: bash tools/gpu_ncu.sh <tag> [reads]
TAG=$1; READS=${2:-1480}
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --reads ${READS} --no-cpu --no-expectations > gpurun_out/${TAG}_bench_ncu_cfg.json 2> gpurun_out/${TAG}_ncu.err && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_cta_kernel<\(int\)2, \(int\)2, \(bool\)0, \(bool\)0>" -c 1 -f -o gpurun_out/${TAG}_cta_full \
    python bench.py --steps 1 --warmup 0 --reads ${READS} --no-cpu --no-expectations > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log | cut -c1-200
