#!/bin/bash
# round 2: ncu captures of the wide kernel (contract pattern, and forced onto the dense pattern)
mkdir -p gpurun_out
python tools/wide_quick.py contract 592 > gpurun_out/r2b_quick_contract.txt 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)0, \(bool\)0>" -s 1 -c 1 -f -o gpurun_out/r2b_wide2048 \
    python tools/wide_quick.py contract 592 > gpurun_out/r2b_ncu1.log 2>&1
tail -2 gpurun_out/r2b_ncu1.log | cut -c1-300
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)0, \(bool\)0>" -c 1 -f -o gpurun_out/r2b_wide256_dense \
    python tools/wide_quick.py dense 1184 SA_WIDE_ONLY=1 > gpurun_out/r2b_ncu2.log 2>&1
tail -2 gpurun_out/r2b_ncu2.log | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2b_launches.csv python tools/wide_quick.py contract 592 > gpurun_out/r2b_ncu3.log 2>&1
