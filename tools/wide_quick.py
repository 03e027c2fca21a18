#!/usr/bin/env python
"""resident-rate probe of the band kernels on one anchor pattern.
usage: wide_quick.py <pattern> <reads> [env settings as K=V ...]; prints GCUPS of sa_batch_run (device-timed)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for kv in sys.argv[3:]:
    k, v = kv.split("=", 1)
    os.environ[k] = v
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
pattern, n = sys.argv[1], int(sys.argv[2])
sa.gpu_init(0)
t0 = time.time()
reads_d, table = synth.make_batch(n, seed=1, **synth.ANCHOR_PATTERNS[pattern])
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
t1 = time.time()
b = sa.Batch(model, reads, p)
t2 = time.time()
cells = b.band_cells
out = dict(pattern=pattern, reads=n, cells=cells, kernel_items=b.kernel_items, wide_items=b.wide_items, gen_s=t1 - t0, prepare_s=t2 - t1,
           env={k: os.environ[k] for k in os.environ if k.startswith("SA_")})
ms = []
for i in range(3):
    b.run()
    m, _ = b.timing()
    ms.append(sum(m))
    out.setdefault('ms_split', []).append([round(v, 1) for v in m])
res = b.fetch()
out["ms"] = ms
out["gcups"] = cells / (min(ms) * 1e-3) / 1e9
out["pairs"] = int(res.pair_offsets[-1])
out["events"] = res.events
print(out, flush=True)
