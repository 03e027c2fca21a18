python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 1 --warmup 1 --reads 4736 --no-cpu > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; tail -3 gpurun_out/bench_f.err; python -c "
import json; d=json.load(open('gpurun_out/bench_f.json')); print(d['ms_per_step'], d['gcups'], d['value'], d['e2e']['gcups'])"
