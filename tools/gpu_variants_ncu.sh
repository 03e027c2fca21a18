#!/bin/bash
# ncu --set full of the EX (expectations) and HD (HDP emission) variants of the band kernel.  usage: bash tools/gpu_variants_ncu.sh <tag>
TAG=$1
mkdir -p gpurun_out
python tools/exp_quick.py 1480 > gpurun_out/${TAG}_exp_quick.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_cta_kernel<\(int\)2, \(int\)2, \(bool\)1, \(bool\)0>" -c 1 -f -o gpurun_out/${TAG}_cta_ex \
    python tools/exp_quick.py 1480 > gpurun_out/${TAG}_ncu_ex.log 2>&1
python tools/hdp_quick.py 1480 > gpurun_out/${TAG}_hdp_quick.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_cta_kernel<\(int\)2, \(int\)2, \(bool\)0, \(bool\)1>" -c 1 -f -o gpurun_out/${TAG}_cta_hd \
    python tools/hdp_quick.py 1480 > gpurun_out/${TAG}_ncu_hd.log 2>&1
tail -3 gpurun_out/${TAG}_exp_quick.log gpurun_out/${TAG}_hdp_quick.log; tail -2 gpurun_out/${TAG}_ncu_ex.log gpurun_out/${TAG}_ncu_hd.log | cut -c1-200
