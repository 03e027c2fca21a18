python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 2 --warmup 1 --reads 1184 --no-cpu > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; tail -3 gpurun_out/bench_f.err; python -c "
import json; d=json.load(open('gpurun_out/bench_f.json')); print(d['ms_per_step'], d['gcups'], d['value'], d['e2e'], d['config']['work_items_by_kernel_rank0'])"
