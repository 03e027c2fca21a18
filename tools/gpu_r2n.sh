#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2n}
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide32_kernel<\(bool\)1, \(int\)2048" -c 1 -f -o gpurun_out/${T}_wide32_2048 \
    python tools/wide_quick.py contract 1184 SA_PRECISION=fp32 > gpurun_out/${T}_ncu1.log 2>&1
tail -1 gpurun_out/${T}_ncu1.log | cut -c1-200
ls -la gpurun_out/
