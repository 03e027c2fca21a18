python bench.py --steps 1 --warmup 1 --reads 592 --no-cpu > gpurun_out/b4.json 2> gpurun_out/b4.err && \
ncu --set full --clock-control none --import-source on -k regex:sa_fast_kernel -c 1 -f -o gpurun_out/prof_fast_r1 python bench.py --steps 1 --warmup 0 --reads 592 --no-cpu > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu4.log | cut -c1-200
