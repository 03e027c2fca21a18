#!/usr/bin/env python
"""Where the end-to-end time of one batch goes (host pack / H2D / kernels / compaction / D2H), SA_TIMING=1 on stderr."""
import os, sys, time
os.environ["SA_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
sa.gpu_init(0)
reads_d, table = synth.make_batch(n, seed=1, **bench.GEN)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
for i in range(3):
    print("---- pass", i, file=sys.stderr)
    t0 = time.time(); b = sa.Batch(model, reads, p)
    t1 = time.time(); b.run()
    t2 = time.time(); r = b.fetch()
    t3 = time.time(); b.close()
    t4 = time.time()
    print("[py] prepare %.1f ms, run %.1f ms, fetch %.1f ms, close %.1f ms, total %.1f ms" %
          tuple(1e3 * x for x in (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t4 - t0)), file=sys.stderr)
for nch in (1, 2, 4, 8):
    os.environ["SA_CHUNKS"] = str(nch)
    for i in range(3):
        print("---- one-shot, SA_CHUNKS=%d, pass %d" % (nch, i), file=sys.stderr)
        t0 = time.time(); r = sa.align_batch(model, reads, p); t1 = time.time()
        print("[py] align_batch %.1f ms (%d pairs)" % (1e3 * (t1 - t0), int(r.pair_offsets[-1])), file=sys.stderr)
        del r
