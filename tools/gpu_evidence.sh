#!/bin/bash
# round-2 evidence run: bench lines (no profiler), then the ncu launch list of the bench command, DRAM bytes per cell in
# both precisions, and one --set full capture of the dominant kernel in each precision
mkdir -p gpurun_out
T=${1:-evidence}
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
timeout 900 python bench.py --workload config4 --steps 3 --warmup 3 > gpurun_out/${T}_bench_config4.json 2> gpurun_out/${T}_bench_config4.err
timeout 1200 python bench.py --workload config5 --steps 2 --warmup 1 > gpurun_out/${T}_bench_config5.json 2> gpurun_out/${T}_bench_config5.err
timeout 900 python bench.py --precision fp64 --no-other-precision --no-dense --no-expectations --steps 3 --warmup 3 > gpurun_out/${T}_bench_fp64.json 2> gpurun_out/${T}_bench_fp64.err
timeout 900 python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_bench_s2w1.json 2> gpurun_out/${T}_bench_s2w1.err
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__grid_size,launch__block_size --clock-control none --kernel-name-base demangled -c 600 --csv \
    --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_ncu_bench.log 2>&1
for P in fp32 fp64; do
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,launch__grid_size,launch__block_size --clock-control none --kernel-name-base demangled -c 200 --csv \
    --log-file gpurun_out/${T}_dram_${P}.csv python tools/wide_quick.py contract 1184 SA_PRECISION=$P > gpurun_out/${T}_dram_${P}.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide32_kernel<\(bool\)1, \(int\)2048" -c 1 -f -o gpurun_out/${T}_wide32_2048 \
    python tools/wide_quick.py contract 1184 SA_PRECISION=fp32 > gpurun_out/${T}_ncu_w32.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:sa_wide_kernel<\(bool\)0, \(bool\)0, \(bool\)1, \(int\)2048" -c 1 -f -o gpurun_out/${T}_wide64_2048 \
    python tools/wide_quick.py contract 1184 > gpurun_out/${T}_ncu_w64.log 2>&1
ls -la gpurun_out | tail -20
