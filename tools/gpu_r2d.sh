#!/bin/bash
mkdir -p gpurun_out
T=${1:-r2d}
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)0, \(bool\)0>" -c 1 -f -o gpurun_out/${T}_wide2048 \
    python tools/wide_quick.py contract 592 > gpurun_out/${T}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)0, \(bool\)0>" -c 1 -f -o gpurun_out/${T}_wide256_dense \
    python tools/wide_quick.py dense 1184 SA_WIDE_ONLY=1 > gpurun_out/${T}_ncu2.log 2>&1
