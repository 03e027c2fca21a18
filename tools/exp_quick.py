#!/usr/bin/env python
"""throughput of the expectations call (config 4 shape) on n reads"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
n = int(sys.argv[1]) if len(sys.argv) > 1 else 592
sa.gpu_init(0)
reads_d, table = synth.make_batch(n, seed=1, **bench.GEN)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
for i in range(3):
    t0 = time.time(); r = sa.expectations_batch(model, reads, p); t1 = time.time()
    cells = r.band_cells
    print("expectations: %d reads, %.3e cells, %.1f ms, %.2f GCUPS, %.3e events/s; likelihood sum %.6f" %
          (n, cells, 1e3 * (t1 - t0), cells / (t1 - t0) / 1e9, r.events / (t1 - t0), float(r.likelihood.sum())))
b = sa.Batch(model, reads, p, expectations=True)
for i in range(2):
    b.run()
    ms, _ = b.timing()
    print("expectations resident: kernel %.1f ms, %.2f GCUPS" % (sum(ms), b.band_cells / sum(ms) / 1e6))
