"""-m gpu, collected LAST on purpose: sa_gpu_align_noalloc, the allocation-free entry
point (one read through the wide band kernel with every device buffer in a static 16 MiB device array; no counterpart in the
reference), called through the C-ABI and compared with the existing entry
point sa_gpu_align_batch on the same reads — on several calls in a row, because every call shares the static array and
finds the previous call's contents in it."""
import numpy as np
import pytest

import signalalign_b200 as sa
from signalalign_b200 import synth

from helpers import compare_pairs, sa_read

pytestmark = pytest.mark.gpu


def _setup():
    sa.gpu_init(0)
    dense, table = synth.make_batch(3, seed=41, mean_len=700, min_len=300, max_len=1200, trim=0)
    gapped, _ = synth.make_batch(2, seed=42, mean_len=1500, min_len=900, max_len=2200, table=table,
                                 **synth.ANCHOR_PATTERNS["long_runs_trim14"])
    model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
    return dense + gapped, model


def test_noalloc_entry_point_matches_align_batch_on_calls_in_a_row():
    reads, model = _setup()
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    rs = [sa_read(r) for r in reads]
    want = sa.align_batch(model, rs, p)            # the existing entry point (it also puts the model on the device)
    # a long read, then shorter ones, then the first again: stale rows, headers and pairs of earlier calls stay in the array
    order = [4, 0, 3, 1, 2, 4, 0]
    for i in order:
        got = sa.align_noalloc(model, rs[i], p)
        ref = want.read_pairs(i)
        assert len(got) == len(ref) and len(got) > 0, (i, len(got), len(ref))
        assert compare_pairs(got, ref.astype(np.int64), p.threshold) <= 20, i        # same order, same integers
    # other parameters: many traceback blocks, ragged ends off
    p2 = sa.default_params(threshold=0.05, expansion=50, trace_back=20, min_diags=150)
    rs2 = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"], False, False) for r in reads[:2]]
    want2 = sa.align_batch(model, rs2, p2)
    for i in (0, 1, 0):
        got = sa.align_noalloc(model, rs2[i], p2)
        assert compare_pairs(got, want2.read_pairs(i).astype(np.int64), p2.threshold) <= 20, i


def test_noalloc_entry_point_rejects_what_does_not_fit():
    reads, model = _setup()
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    sa.align_batch(model, [sa_read(reads[0])], p)
    r = reads[0]
    big = sa.Read(r["target"], np.zeros((40000, 4)), np.zeros((0, 2)))
    with pytest.raises(sa.SaError, match="too large for the static scratch"):
        sa.align_noalloc(model, big, p)
    with pytest.raises(sa.SaError, match="alphabet"):
        sa.align_noalloc(model, sa.Read(r["target"][:50] + "X" + r["target"][51:], r["events"], r["anchors"]), p)
    with pytest.raises(sa.SaError, match="pairs"):                     # caller's buffer too small: count reported, nothing written
        sa.align_noalloc(model, sa_read(r), p, pair_cap=10)
    fresh = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), synth.gaussian_table(seed=9))
    with pytest.raises(sa.SaError, match="not resident"):              # no upload here: that would allocate
        sa.align_noalloc(fresh, sa_read(r), p)
    # and the array is still good for the next call
    want = sa.align_batch(model, [sa_read(r)], p)
    assert compare_pairs(sa.align_noalloc(model, sa_read(r), p), want.read_pairs(0).astype(np.int64), p.threshold) <= 20
