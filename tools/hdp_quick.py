#!/usr/bin/env python
"""throughput of the alignment call with the HDP emission (config 5's emission family) on n config-3-shaped reads"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2368
G = int(sys.argv[2]) if len(sys.argv) > 2 else 1200
sa.gpu_init(0)
reads_d, table = synth.make_batch(n, seed=1, **bench.GEN)
rng = np.random.default_rng(3)
K = table.shape[0]
grid = np.linspace(50.0, 140.0, G)
mu = table[:, 0]
pp = 0.7 * np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / 2.0) ** 2) / (2.0 * np.sqrt(2 * np.pi)) + \
     0.3 * np.exp(-0.5 * ((grid[None, :] - mu[:, None] - 3.0) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
sl = np.gradient(pp, grid, axis=1)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table, kind=1)
model.set_hdp(grid, np.arange(K, dtype=np.int32), pp, sl)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
b = sa.Batch(model, reads, p)
print("kernel items (general, 32/64, 128/256):", b.kernel_items)
for i in range(3):
    b.run()
    ms, _ = b.timing()
    print("HDP G=%d K=%d rows: %d reads, %.3e cells, kernel %.1f ms, %.2f GCUPS" % (G, K, n, b.band_cells, sum(ms), b.band_cells / sum(ms) / 1e6))
r = b.fetch()
print("pairs", int(r.pair_offsets[-1]))
if os.environ.get("HDP_EXPECT"):
    for i in range(3):
        t0 = time.time(); e = sa.expectations_batch(model, reads, p); dt = time.time() - t0
        print("HDP expectations end to end: %.1f ms, %.2f GCUPS, %d assignments" % (1e3 * dt, b.band_cells / dt / 1e9, int(e.assign_offsets[-1])))
