#!/usr/bin/env python
"""one-shot align_batch timing for a few relative chunk sizes (SA_CHUNK_FRACS); SA_TIMING timeline on stderr"""
import os, sys, time
os.environ["SA_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
fracs = sys.argv[2:] or ["1,1,1,1"]
sa.gpu_init(0)
reads_d, table = synth.make_batch(n, seed=1, **bench.GEN)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
r = sa.align_batch(model, reads, p); del r
for fr in fracs:
    os.environ["SA_CHUNK_FRACS"] = fr
    best = 1e9
    for i in range(3):
        print("---- SA_CHUNK_FRACS=%s pass %d" % (fr, i), file=sys.stderr)
        t0 = time.time(); r = sa.align_batch(model, reads, p); t1 = time.time()
        best = min(best, 1e3 * (t1 - t0))
        del r
    print("fracs %-24s best of 3: %.1f ms" % (fr, best))
