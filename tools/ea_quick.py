#!/usr/bin/env python
"""throughput probe of sa_gpu_event_align_batch: n synthetic reads of ~10 k events, wall clock of the call
(host buffers in, host result out); optionally the numpy oracle on one read for scale.  usage: ea_quick.py [reads]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import signalalign_b200 as sa
from signalalign_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1184
sa.gpu_init(0)
rng = np.random.default_rng(5)
table = synth.gaussian_table("ACGT", 6)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
ref_codes = rng.integers(0, 4, 2_000_000).astype(np.int64)
reads = []
for _ in range(n):
    r = synth.make_read(rng, ref_codes, table, "ACGT", 6, n_probs=(0.0, 0.70, 0.22, 0.08))   # (every kmer emits: a random table
    # makes a skipped kmer cost ~50 nats, which alone fails the reference's -5.2 average-emission QC)
    reads.append(dict(sequence=r["target"], event_mean=r["events"][:, 0].astype(np.float32), rna=False,
                      scalings=(r["scale"], r["shift"], r["var"])))
ev = sum(len(r["event_mean"]) for r in reads)
cells = sum((len(r["event_mean"]) + len(r["sequence"]) - 4) * 100 for r in reads)
sa.event_align_batch(model, reads[:8])
best = 1e9
for _ in range(3):
    t0 = time.time(); res = sa.event_align_batch(model, reads); best = min(best, time.time() - t0)
print(dict(reads=n, events=ev, failed=sum(r["failed"] for r in res), call_ms=1e3 * best, events_per_s=ev / best, band_cells_per_s=cells / best))
