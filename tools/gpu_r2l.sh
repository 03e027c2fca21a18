#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2l}
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide32_kernel<\(bool\)1>" -c 1 -f -o gpurun_out/${T}_wide32_gr \
    python tools/wide_quick.py contract 592 SA_PRECISION=fp32 SA_WIDE_NT=128,256,256,256,512 > gpurun_out/${T}_ncu1.log 2>&1
tail -1 gpurun_out/${T}_ncu1.log | cut -c1-200
