"""Size-independent properties of the alignment at the benchmark's batch shape (the oracle needs ~1.5 s of CPU per read at
this size, so the full batch is checked through invariants instead; single reads of this size are compared with the
oracle in tests/test_gpu_wide.py and tests/test_gpu_wide32.py):

  * determinism: two calls on the same batch return byte-identical results (work items are pulled from an atomic queue
    in a run-dependent order; nothing may depend on it);
  * the staged batch and the pipelined one-shot call agree byte for byte;
  * coordinates lie inside each read, every posterior is in [threshold, 1], the kmer of a pair is the reference's kmer at x;
  * posterior mass: the match posteriors of one event over all reference positions sum to at most 1 (up to the reference's
    logAdd approximation, 5.5e-4 per term);
  * both precisions: same pair set up to the threshold band, posteriors within the measured FP32 envelope."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

pytestmark = pytest.mark.gpu
N_READS = 296


def _batch():
    import bench
    import signalalign_b200 as sa
    from signalalign_b200 import synth
    reads_d, table = synth.make_batch(N_READS, seed=99, **bench.GEN)
    model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
    p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
    reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
    return sa, synth, reads_d, table, model, p, reads


def test_benchmark_shape_batch_invariants():
    sa, synth, reads_d, table, model, p, reads = _batch()
    sa.gpu_init(0)
    sa.set_precision("fp64")
    b = sa.Batch(model, reads, p)
    assert b.kernel_items[0] == 0, "no read of the contract pattern may fall to the general kernel"
    b.run()
    r1 = b.fetch()
    pairs1, offs1 = r1.pairs.copy(), r1.pair_offsets.copy()
    b.run()
    r2 = b.fetch()
    assert np.array_equal(offs1, r2.pair_offsets) and np.array_equal(pairs1, r2.pairs), "two runs differ"
    b.close()
    r3 = sa.align_batch(model, reads, p)
    assert np.array_equal(offs1, r3.pair_offsets) and np.array_equal(pairs1, r3.pairs), "staged batch and one-shot call differ"
    thr = int(np.floor(0.01 * 1e7))
    for i, rd in enumerate(reads_d):
        pr = pairs1[offs1[i]:offs1[i + 1]]
        lx, ly = len(rd["target"]) - 5, rd["events"].shape[0]
        assert len(pr) > 0.5 * ly
        assert pr[:, 1].min() >= 0 and pr[:, 1].max() < lx and pr[:, 2].min() >= 0 and pr[:, 2].max() < ly
        assert pr[:, 0].min() >= thr and pr[:, 0].max() <= 10_000_000
        codes = np.frombuffer(rd["target"].encode(), np.uint8)
        lut = np.zeros(256, np.int64)
        for j, ch in enumerate("ACGT"):
            lut[ord(ch)] = j
        kid = synth.kmer_indices(lut[codes], 6, 4)
        assert np.array_equal(pr[:, 3], kid[pr[:, 1]])
        mass = np.bincount(pr[:, 2], weights=pr[:, 0] / 1e7, minlength=ly)
        assert mass.max() <= 1.0 + 5e-3, mass.max()
    # the FP32 mode on the same batch: same pairs away from the threshold, posteriors inside its measured envelope
    sa.set_precision("fp32")
    try:
        r4 = sa.align_batch(model, reads, p)
    finally:
        sa.set_precision("fp64")
    import bench
    agree = bench.compare_pairs(r3, r4)
    assert agree["kmer_mismatches"] == 0
    assert agree["max_abs_dp"] <= 6e-4 and agree["dp_above_1e-4"] <= 1e-5 * agree["pairs_compared"]
    assert agree["pairs_in_one_only"] <= 1e-5 * agree["pairs_compared"] and (agree["largest_p_of_those"] or 0.0) <= 0.0105
