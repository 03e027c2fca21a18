#!/usr/bin/env python
"""BASELINE config 2's shape at throughput size: cytosine-variant calling.  ACEGT 6-mer model (E = 5mC: the C kmer's
level shifted), config-3-shaped ACGT reads, a fraction of the reference's C positions replaced by the two-way ambiguity
code P (C/E, create_ambig_bases: impl/pairwiseAligner.c:32-65) -> multi-path cells -> the general kernel.
usage: tools/config2.py [reads] [fraction of C flagged] [hdp grid length, 0 = Gaussian]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from signalalign_b200 import synth
import signalalign_b200 as sa
n = int(sys.argv[1]) if len(sys.argv) > 1 else 592
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.1
G = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sa.gpu_init(0)
reads_d, t4 = synth.make_batch(n, seed=1, **bench.GEN)
# ACEGT table from the ACGT one: every kmer takes the row of its de-methylated twin, + 1.5 pA per E
A5 = "ACEGT"
K5 = 5 ** 6
digits = np.zeros((K5, 6), np.int64)
v = np.arange(K5)
for j in range(5, -1, -1):
    digits[:, j] = v % 5
    v //= 5
to4 = np.array([0, 1, 1, 2, 3])          # A C E G T -> A C C G T
twin = np.zeros(K5, np.int64)
for j in range(6):
    twin = twin * 4 + to4[digits[:, j]]
t5 = t4[twin].copy()
t5[:, 0] += 1.5 * (digits == 2).sum(1)
model = sa.Model.from_tables(A5, 6, synth.nanopore_transitions(), t5, kind=1 if G else 0)
if G:
    grid = np.linspace(50.0, 140.0, G)
    mu = t5[:, 0]
    pp = 0.7 * np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / 2.0) ** 2) / (2.0 * np.sqrt(2 * np.pi)) + \
         0.3 * np.exp(-0.5 * ((grid[None, :] - mu[:, None] - 3.0) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
    model.set_hdp(grid, np.arange(K5, dtype=np.int32), pp, np.gradient(pp, grid, axis=1))
rng = np.random.default_rng(11)
reads = []
nflag = 0
for r in reads_d:
    t = np.frombuffer(r["target"].encode(), np.uint8).copy()
    c = np.nonzero(t == ord("C"))[0]
    pick = c[rng.random(len(c)) < frac]
    t[pick] = ord("P")
    nflag += len(pick)
    reads.append(sa.Read(t.tobytes().decode(), r["events"], r["anchors"], r["scale"], r["shift"], r["var"]))
p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
b = sa.Batch(model, reads, p)
print("kernel items (general, 32/64, 128/256):", b.kernel_items, "flagged positions per read: %.1f" % (nflag / n))
for i in range(2):
    b.run()
    ms, _ = b.timing()
    print("config-2 shape %s: %d reads, %.3e band cells, kernel %.1f ms, %.2f GCUPS (band cells)" %
          ("HDP G=%d" % G if G else "Gaussian", n, b.band_cells, sum(ms), b.band_cells / sum(ms) / 1e6))
r = b.fetch()
print("pairs", int(r.pair_offsets[-1]), "path cells %.3e" % r.path_cells)
