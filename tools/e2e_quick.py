#!/usr/bin/env python
"""one-shot align_batch timing for a few SA_CHUNKS / env settings (SA_TIMING on stderr)"""
import os, sys, time
os.environ["SA_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
chunks = [int(c) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 4]
sa.gpu_init(0)
reads_d, table = synth.make_batch(n, seed=1, **bench.GEN)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
for nch in chunks:
    os.environ["SA_CHUNKS"] = str(nch)
    for i in range(3):
        print("---- one-shot, SA_CHUNKS=%d, pass %d" % (nch, i), file=sys.stderr)
        t0 = time.time(); r = sa.align_batch(model, reads, p); t1 = time.time()
        print("[py] align_batch %.1f ms (%d pairs)" % (1e3 * (t1 - t0), int(r.pair_offsets[-1])), file=sys.stderr)
        del r
