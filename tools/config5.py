#!/usr/bin/env python
"""BASELINE config 5's shape: RNA 5-mer table (K = 1024), HDP emission (synthetic spline tables: the RNA .nhdp is absent
from the reference snapshot), read lengths log-normal with a long tail (median ~20 k events, sigma 1.35, clipped to
[1000, 400000] events), un-ambiguous reference.  Reports resident and end-to-end throughput of the alignment call and
which kernel classes the length/band bucketing chose."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from signalalign_b200 import synth
import signalalign_b200 as sa

n = int(sys.argv[1]) if len(sys.argv) > 1 else 768
hdp = (sys.argv[2] if len(sys.argv) > 2 else "hdp") == "hdp"
sa.gpu_init(0)
# ~1.3 events per nt for RNA: median 20 k events = 15.4 k nt
n_probs = (0.05, 0.70, 0.18, 0.07)
t0 = time.time()
reads_d, table = synth.make_batch(n, seed=20261105, alphabet="ACGT", k=5, ref_len=3_000_000, mean_len=15400,
                                  sigma_len=1.35, min_len=770, max_len=300000, trim=0, block_mean=80, n_probs=n_probs)
ev = np.array([r["events"].shape[0] for r in reads_d])
print("reads %d: events median %d mean %d max %d (generated in %.1f s)" % (n, np.median(ev), ev.mean(), ev.max(), time.time() - t0))
K = table.shape[0]
kind = 1 if hdp else 0
model = sa.Model.from_tables("ACGT", 5, synth.nanopore_transitions(), table, kind=kind)
if hdp:
    G = 1200
    grid = np.linspace(50.0, 140.0, G)
    mu = table[:, 0]
    pp = 0.7 * np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / 2.0) ** 2) / (2.0 * np.sqrt(2 * np.pi)) + \
         0.3 * np.exp(-0.5 * ((grid[None, :] - mu[:, None] - 3.0) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
    model.set_hdp(grid, np.arange(K, dtype=np.int32), pp, np.gradient(pp, grid, axis=1))
p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
b = sa.Batch(model, reads, p)
print("kernel items (general, 32/64 slots, 128/256 slots):", b.kernel_items, "band cells %.3e" % b.band_cells)
for i in range(3):
    b.run()
    ms, _ = b.timing()
    print("%s resident: %.1f ms, %.2f GCUPS, %.3e events/s" % ("HDP" if hdp else "Gaussian", sum(ms), b.band_cells / sum(ms) / 1e6, ev.sum() / sum(ms) * 1e3))
cells = b.band_cells
b.close()
for i in range(0 if os.environ.get("C5_RESIDENT_ONLY") else 2):
    t1 = time.time(); r = sa.align_batch(model, reads, p); dt = time.time() - t1
    print("end to end: %.1f ms, %.2f GCUPS, %.3e events/s, %d pairs" % (1e3 * dt, cells / dt / 1e9, ev.sum() / dt, int(r.pair_offsets[-1])))
    del r
