#!/bin/bash
# A/B of compile-time variants of the band kernel.  Here (no GPU):  bash tools/ab.sh build NAME "-DFLAG=.. -DFLAG2=.."
# builds gpurun_ab/libsagpu_NAME.so; on the GPU box:  bash tools/ab.sh run NAME...  prints one bench line per variant.
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
    mkdir -p gpurun_ab
    make -s -C signalalign_b200/csrc OUT=$PWD/gpurun_ab/libsagpu_$2.so EXTRA="$3"
else
    shift
    for v in "$@"; do
        SA_LIBSAGPU=$PWD/gpurun_ab/libsagpu_$v.so python bench.py --no-cpu --no-expectations --steps 2 > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || tail -3 gpurun_out/ab_$v.err
        python tools/ab_line.py $v gpurun_out/ab_$v.json
    done
fi
