#!/bin/bash
# One GPU-box visit: parity tests, smoke, peak micro-benchmark, both bench arms, ncu launch list, one full ncu capture.
# usage (from the repo root on the box): bash tools/gpu_round.sh <tag> [reads-for-ncu]
TAG=${1:-r1}
NCU_READS=${2:-2960}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
nproc > gpurun_out/${TAG}_nproc.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.txt 2>&1
[ -x tools/peak_fp ] || nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/peak_fp tools/peak_fp.cu
tools/peak_fp > gpurun_out/${TAG}_peak_fp.jsonl 2>&1
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?" >> gpurun_out/${TAG}_bench.err
python bench.py --steps 2 --warmup 1 --no-cpu --no-expectations > gpurun_out/${TAG}_bench_short.json 2>> gpurun_out/${TAG}_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-expectations > gpurun_out/${TAG}_ncu_launches.log 2>&1
python bench.py --steps 1 --warmup 1 --reads ${NCU_READS} --no-cpu --no-expectations > gpurun_out/${TAG}_bench_ncu_cfg.json 2>> gpurun_out/${TAG}_bench.err && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_cta_kernel<\(int\)2, \(int\)2, \(bool\)0, \(bool\)0>" -c 1 -f -o gpurun_out/${TAG}_cta_full \
    python bench.py --steps 1 --warmup 0 --reads ${NCU_READS} --no-cpu --no-expectations > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -3 gpurun_out/${TAG}_pytest.txt; cat gpurun_out/${TAG}_smoke.txt | tail -2; cat gpurun_out/${TAG}_peak_fp.jsonl
cut -c1-600 gpurun_out/${TAG}_bench_reference.json; cut -c1-2500 gpurun_out/${TAG}_bench.json
