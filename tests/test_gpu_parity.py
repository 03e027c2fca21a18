"""-m gpu: the CUDA path, called through the C-ABI, against the oracle on the same seeded inputs."""
import os

import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port
from signalalign_b200 import synth

from helpers import (compare_expectations, compare_pairs, oracle_align, oracle_expectations, port_model_from, port_params,
                     sa_read)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small_batch():
    # dense guide anchors (trim 0): bands stay ~expansion+1 wide -> fast kernel
    reads, table = synth.make_batch(6, seed=11, mean_len=500, min_len=200, max_len=900, trim=0)
    return reads, table


@pytest.fixture(scope="module")
def gapped_batch():
    # anchors trimmed by 14 per block (constraintDiagonalTrim): bands widen in the gaps -> mixed kernels
    reads, table = synth.make_batch(6, seed=12, mean_len=500, min_len=200, max_len=900, trim=14)
    return reads, table


def _run_and_compare(reads, table, p, max_dp=10, force_general=False, expect_fast=None, no_latency=False):
    """no_latency: keep small batches on the throughput mapping (two cells per thread); by default the scheduler sends
    reads that dominate their batch to the one-cell mapping, which is every read of a six-read test batch."""
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    pm = port_model_from("ACGT", 6, trans, table)
    if force_general:
        os.environ["SA_FORCE_GENERAL"] = "1"
    if no_latency:
        os.environ["SA_NO_LATENCY_CLASS"] = "1"
    try:
        b = sa.Batch(model, [sa_read(r) for r in reads], p)
        kernels = b.kernel_items
        b.run()
        res = b.fetch()
        b.close()
    finally:
        os.environ.pop("SA_FORCE_GENERAL", None)
        os.environ.pop("SA_NO_LATENCY_CLASS", None)
    if force_general:
        assert kernels[1] == 0 and kernels[2] == 0
    elif expect_fast is not None:
        assert kernels[1] + kernels[2] == len(reads), kernels
    worst = 0
    cells = 0
    for i, r in enumerate(reads):
        ora, stats = oracle_align(pm, r, port_params(p))
        cells += stats[0]
        worst = max(worst, compare_pairs(res.read_pairs(i), ora, p.threshold))
    assert res.band_cells == cells  # same integer band logic as the reference's band_construct
    assert worst <= max_dp
    return worst


@pytest.mark.parametrize("force_general", [False, True])
@pytest.mark.parametrize("expansion,trace_back,min_diags", [(50, 50, 200), (100, 50, 300), (20, 40, 1000), (6, 2, 10)])
def test_gaussian_meanonly_small(small_batch, expansion, trace_back, min_diags, force_general):
    reads, table = small_batch
    p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
    expect = {50: 1, 100: 2, 20: 1, 6: 1}[expansion]
    _run_and_compare(reads, table, p, force_general=force_general, expect_fast=expect)
    if not force_general and expansion == 100:      # 128-slot class: both mappings
        _run_and_compare(reads, table, p, expect_fast=expect, no_latency=True)


@pytest.mark.parametrize("expansion", [20, 50, 100])
def test_gapped_anchors_mixed_kernels(gapped_batch, expansion):
    reads, table = gapped_batch
    p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=20, min_diags=150)
    _run_and_compare(reads, table, p)


def test_ragged_flags_and_thresholds(small_batch):
    reads, table = small_batch
    for rl, rr, thr in [(False, False, 0.01), (True, False, 0.2), (False, True, 0.001)]:
        rs = [dict(r, ragged_left=rl, ragged_right=rr) for r in reads[:3]]
        p = sa.default_params(threshold=thr, expansion=50, trace_back=50, min_diags=200)
        _run_and_compare(rs, table, p)


def test_pair_slots_overflow_and_grow_in_every_mode(small_batch):
    """threshold 1e-7: far more than 3 pairs per event, so the first slot size overflows and the batch is re-slotted and
    re-run — in the latency class (every read of a six-read batch), on the throughput mapping and on the general kernel"""
    reads, table = small_batch
    p = sa.default_params(threshold=1e-7, expansion=100, trace_back=50, min_diags=300)
    _run_and_compare(reads[:3], table, p)
    _run_and_compare(reads[:3], table, p, no_latency=True)
    _run_and_compare(reads[:3], table, p, force_general=True)


def test_medium_reads_many_blocks():
    reads, table = synth.make_batch(3, seed=5, mean_len=2500, min_len=2000, max_len=3000, trim=0)
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    _run_and_compare(reads, table, p, expect_fast=2)
    _run_and_compare(reads, table, p, expect_fast=2, no_latency=True)


@pytest.mark.parametrize("expansion,trace_back,min_diags", [(50, 50, 200), (100, 50, 300), (20, 40, 1000), (6, 2, 10)])
def test_expectations_band_kernel_vs_oracle_and_general(small_batch, gapped_batch, expansion, trace_back, min_diags):
    """getExpectationsUsingAnchors on the band kernel (<W, C, EX = true>): transitions and likelihood against the oracle
    (1e-6 / 1e-5 relative) and against the general kernel on the same batch; ragged flags on and off."""
    trans = synth.nanopore_transitions()
    for reads, table in (small_batch, gapped_batch):
        model = sa.Model.from_tables("ACGT", 6, trans, table)
        pm = port_model_from("ACGT", 6, trans, table)
        p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
        rs = [dict(r, ragged_left=(i % 2 == 0), ragged_right=(i % 3 != 0)) for i, r in enumerate(reads)]
        b = sa.Batch(model, [sa_read(r) for r in rs], p, expectations=True)
        kernels = b.kernel_items
        b.run()
        fast = b.fetch()
        b.close()
        if reads is small_batch[0]:
            assert kernels[1] + kernels[2] == len(rs), kernels     # dense anchors: every read on the band kernel
        os.environ["SA_FORCE_GENERAL"] = "1"
        try:
            gen = sa.expectations_batch(model, [sa_read(r) for r in rs], p)
        finally:
            os.environ.pop("SA_FORCE_GENERAL", None)
        for i, r in enumerate(rs):
            t, lik = oracle_expectations(pm, r, port_params(p))[:2]
            compare_expectations(fast.transitions[i], fast.likelihood[i], t, lik)
            compare_expectations(fast.transitions[i], fast.likelihood[i], gen.transitions[i], gen.likelihood[i])


def test_device_logadd_bit_exact():
    """The device logAdd against the oracle's (itself bit-identical to the compiled reference) on random pairs,
    the range seams (1.0, 2.5, 4.5), the 7.5 cut-off and -inf operands."""
    rng = np.random.default_rng(3)
    n = 200000
    x = rng.uniform(-50, 5, n)
    y = x + rng.uniform(-9, 9, n)
    seams = np.array([0.0, 1.0, 2.5, 4.5, 7.5])
    eps = np.array([0.0, 1e-15, -1e-15, 2.3e-16, 1e-9, -1e-9, 4.4e-16, 8.9e-16])
    sx = np.repeat(rng.uniform(-30, 0, len(seams) * len(eps)), 2)
    d = (seams[:, None] + eps[None, :]).reshape(-1)
    sy = sx.copy()
    sy[0::2] = sx[0::2] + d
    sy[1::2] = sx[1::2] - d
    x = np.concatenate([x, sx, [-np.inf, -np.inf, 3.0, -np.inf]])
    y = np.concatenate([y, sy, [-np.inf, 2.0, -np.inf, -1e300]])
    got = sa.debug_logadd(x, y)
    exp = np.array([port.log_add(a, b) for a, b in zip(x, y)])
    assert np.array_equal(got.view(np.int64), exp.view(np.int64))


# ------------------------------------------------------------------------------------------------ every emission kind
def _synthetic_hdp(table, seed=3, G=120):
    """single-level HDP tables in the shape of a .nhdp: a two-component mixture around the table mean per kmer row"""
    rng = np.random.default_rng(seed)
    K = table.shape[0]
    rows = 300
    grid = np.linspace(50.0, 135.0, G)
    centers = rng.uniform(60, 125, rows)
    pp = np.zeros((rows, G))
    for r in range(rows):
        a, b = centers[r], centers[r] + rng.uniform(-6, 6)
        pp[r] = 0.7 * np.exp(-0.5 * ((grid - a) / 2.5) ** 2) / (2.5 * np.sqrt(2 * np.pi)) + \
            0.3 * np.exp(-0.5 * ((grid - b) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
    sl = np.gradient(pp, grid, axis=1)
    # each kmer uses the row whose centre is nearest to its table mean (some rows shared = unobserved kmers' parents)
    k2r = np.abs(table[:, 0][:, None] - centers[None, :]).argmin(1).astype(np.int32)
    return grid, k2r, pp, sl


@pytest.mark.parametrize("kind", [port.EMIT_GAUSS_NOISE_SCALED, port.EMIT_GAUSS_NOISE_DESCALE, port.EMIT_HDP])
def test_other_emission_kinds_alignment_and_expectations(gapped_batch, kind):
    reads, table = gapped_batch
    trans = synth.nanopore_transitions()
    kw = {}
    model = sa.Model.from_tables("ACGT", 6, trans, table, kind=kind)
    if kind == port.EMIT_HDP:
        grid, k2r, pp, sl = _synthetic_hdp(table)
        model.set_hdp(grid, k2r, pp, sl)
        kw = dict(grid=grid, kmer_to_row=k2r, post_pred=pp, slopes=sl)
    pm = port_model_from("ACGT", 6, trans, table, kind=kind, **kw)
    p = sa.default_params(threshold=0.01, expansion=20, trace_back=20, min_diags=150)
    rs = [dict(r) for r in reads[:4]]
    for r in rs:   # realistic event noise for the inverse-Gaussian term
        r["events"] = r["events"].copy()
        r["events"][:, 1] = np.random.default_rng(1).uniform(0.5, 2.5, len(r["events"]))
    res = sa.align_batch(model, [sa_read(r) for r in rs], p)
    exp = sa.expectations_batch(model, [sa_read(r) for r in rs], p)
    for i, r in enumerate(rs):
        ora, _ = oracle_align(pm, r, port_params(p))
        assert compare_pairs(res.read_pairs(i), ora, p.threshold) <= 20
        tr, lik, ak, ae = oracle_expectations(pm, r, port_params(p))
        compare_expectations(exp.transitions[i], exp.likelihood[i], tr, lik)
        if kind == port.EMIT_HDP:
            a0, a1 = exp.assign_offsets[i], exp.assign_offsets[i + 1]
            assert np.array_equal(exp.assign_kmer_pos[a0:a1], ak) and np.array_equal(exp.assign_event[a0:a1], ae)


@pytest.mark.parametrize("expansion,trace_back,min_diags", [(50, 50, 200), (100, 50, 300), (20, 40, 1000)])
def test_hdp_emission_on_the_band_kernel(small_batch, expansion, trace_back, min_diags):
    """Un-ambiguous reference + HDP machine: the band kernel's HDP variant (<W, C, false, HD = true>) against the oracle
    and against the general kernel; events beyond both ends of the spline grid included (linear extrapolation)."""
    reads, table = small_batch
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table, kind=port.EMIT_HDP)
    grid, k2r, pp, sl = _synthetic_hdp(table)
    model.set_hdp(grid, k2r, pp, sl)
    pm = port_model_from("ACGT", 6, trans, table, kind=port.EMIT_HDP, grid=grid, kmer_to_row=k2r, post_pred=pp, slopes=sl)
    p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
    rs = [dict(r) for r in reads]
    rs[1]["events"] = rs[1]["events"].copy()
    rs[1]["events"][::37, 0] = 20.0      # far left of the grid
    rs[1]["events"][5::41, 0] = 190.0    # far right
    b = sa.Batch(model, [sa_read(r) for r in rs], p)
    kernels = b.kernel_items
    b.run()
    fast = b.fetch()
    b.close()
    assert kernels[1] + kernels[2] == len(rs), kernels
    os.environ["SA_FORCE_GENERAL"] = "1"
    try:
        gen = sa.align_batch(model, [sa_read(r) for r in rs], p)
    finally:
        os.environ.pop("SA_FORCE_GENERAL", None)
    for i, r in enumerate(rs):
        ora, _ = oracle_align(pm, r, port_params(p))
        assert compare_pairs(fast.read_pairs(i), ora, p.threshold) <= 20
        assert compare_pairs(fast.read_pairs(i), gen.read_pairs(i), p.threshold) <= 20
    # expectations + assignments of the HDP machine on the band kernel (<W, C, EX = true, HD = true>): transitions and
    # likelihood against the oracle, the assignment list (order included) bit-exact against oracle and general kernel
    fe = sa.expectations_batch(model, [sa_read(r) for r in rs], p)
    os.environ["SA_FORCE_GENERAL"] = "1"
    try:
        ge = sa.expectations_batch(model, [sa_read(r) for r in rs], p)
    finally:
        os.environ.pop("SA_FORCE_GENERAL", None)
    assert np.array_equal(fe.assign_offsets, ge.assign_offsets)
    assert np.array_equal(fe.assign_kmer_pos, ge.assign_kmer_pos) and np.array_equal(fe.assign_event, ge.assign_event)
    assert fe.assign_offsets[-1] > 0
    for i, r in enumerate(rs):
        tr, lik, ak, ae = oracle_expectations(pm, r, port_params(p))
        compare_expectations(fe.transitions[i], fe.likelihood[i], tr, lik)
        compare_expectations(fe.transitions[i], fe.likelihood[i], ge.transitions[i], ge.likelihood[i])
        a0, a1 = fe.assign_offsets[i], fe.assign_offsets[i + 1]
        assert np.array_equal(fe.assign_kmer_pos[a0:a1], ak) and np.array_equal(fe.assign_event[a0:a1], ae)
    # a batch this small is all latency class -> split mode (forward sweep keeping every row and its log-density plane,
    # traceback blocks as work items); ring mode, the block re-run path and the throughput mapping must give the same bytes
    for env in ({"SA_NO_SPLIT": "1"}, {"SA_SPLIT_PAIRCAP": "8"}, {"SA_NO_LATENCY_CLASS": "1"}):
        os.environ.update(env)
        try:
            o = sa.align_batch(model, [sa_read(r) for r in rs], p)
        finally:
            for k in env:
                os.environ.pop(k, None)
        assert np.array_equal(fast.pair_offsets, o.pair_offsets) and np.array_equal(fast.pairs, o.pairs), env


def test_ambiguous_reference_iupac(gapped_batch):
    reads, table = gapped_batch
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    pm = port_model_from("ACGT", 6, trans, table)
    rng = np.random.default_rng(8)
    rs = []
    for r in reads[:3]:
        t = list(r["target"])
        for i in rng.choice(len(t), 14, replace=False):
            t[i] = "RYSWKMBDHVX"[int(rng.integers(0, 11))]
        rs.append(dict(r, target="".join(t)))
    p = sa.default_params(threshold=0.01, expansion=20, trace_back=20, min_diags=150)
    res = sa.align_batch(model, [sa_read(r) for r in rs], p)
    exp = sa.expectations_batch(model, [sa_read(r) for r in rs], p)
    assert res.path_cells > res.band_cells
    for i, r in enumerate(rs):
        ora, _ = oracle_align(pm, r, port_params(p))
        assert compare_pairs(res.read_pairs(i), ora, p.threshold) <= 20
        tr, lik, _, _ = oracle_expectations(pm, r, port_params(p))
        compare_expectations(exp.transitions[i], exp.likelihood[i], tr, lik)


def test_split_regions_large_gaps():
    """anchor gaps larger than splitMatrixBiggerThanThis cut the read into independently aligned regions with ragged
    internal ends (reference pairwiseAligner.c:1953-2016); pairs come back in the reference's region order"""
    reads, table = synth.make_batch(3, seed=21, mean_len=700, min_len=500, max_len=900, trim=14, block_mean=40)
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    pm = port_model_from("ACGT", 6, trans, table)
    p = sa.default_params(threshold=0.01, expansion=20, trace_back=20, min_diags=150, split_bigger=600)
    res = sa.align_batch(model, [sa_read(r) for r in reads], p)
    for i, r in enumerate(reads):
        assert len(port.split_points(r["anchors"], len(r["target"]) - 5, len(r["events"]), 600, True, True)) > 1
        ora, _ = oracle_align(pm, r, port_params(p))
        assert compare_pairs(res.read_pairs(i), ora, p.threshold) <= 20


def test_rna_5mer_table_long_read():
    """config 5's table shape (ACGT 5-mer, K = 1024) on a longer read"""
    table = synth.gaussian_table("ACGT", 5, seed=11)
    reads, _ = synth.make_batch(2, seed=13, k=5, table=table, mean_len=5000, min_len=4000, max_len=6000, trim=0,
                                n_probs=(0.15, 0.60, 0.18, 0.07))
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 5, trans, table)
    pm = port_model_from("ACGT", 5, trans, table)
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=1000)
    res = sa.align_batch(model, [sa_read(r) for r in reads], p)
    for i, r in enumerate(reads):
        ora, _ = oracle_align(pm, r, port_params(p))
        assert compare_pairs(res.read_pairs(i), ora, p.threshold) <= 20


# ------------------------------------------------------------------------------------------------ full-size properties
def test_config3_size_reads_band_kernel_equals_general_kernel():
    """BASELINE config 3's shape (~10 k events per read, -x 100): the band kernel and the general kernel are two
    independent implementations of the same recurrence (thread-per-cell sliding window vs. CTA-wide diagonals over
    explicit band tables); at full read size they must agree pair for pair (order, coordinates, kmers; posteriors
    within the FP64 bound), and so must their expectations.  The oracle checks three of the reads in full."""
    import bench
    reads, table = synth.make_batch(48, seed=77, **bench.GEN)
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK,
                          min_diags=bench.MIN_DIAGS)
    rs = [sa_read(r) for r in reads]
    # 48 reads: every read dominates the batch -> split mode (forward sweep on the one-cell mapping keeping every row,
    # traceback blocks as independent work items); the same through the ring-mode latency class, through the block
    # kernel's re-run path (slots of 8 pairs), and on the throughput mapping (two cells per thread): all bit-identical
    lat = sa.align_batch(model, rs, p)
    outs = []
    for env in ({"SA_NO_SPLIT": "1"}, {"SA_SPLIT_PAIRCAP": "8"}, {"SA_NO_LATENCY_CLASS": "1"}):
        os.environ.update(env)
        try:
            outs.append(sa.align_batch(model, rs, p))
        finally:
            for k in env:
                os.environ.pop(k, None)
    fast = outs[2]
    for o in outs:
        assert np.array_equal(lat.pair_offsets, o.pair_offsets) and np.array_equal(lat.pairs, o.pairs)
    fe = sa.expectations_batch(model, rs, p)
    os.environ["SA_FORCE_GENERAL"] = "1"
    try:
        gen = sa.align_batch(model, rs, p)
        ge = sa.expectations_batch(model, rs, p)
    finally:
        os.environ.pop("SA_FORCE_GENERAL", None)
    assert fast.band_cells == gen.band_cells and fast.events == gen.events
    assert np.array_equal(fast.pair_offsets, gen.pair_offsets)
    assert np.array_equal(fast.pairs[:, 1:], gen.pairs[:, 1:])
    assert np.abs(fast.pairs[:, 0].astype(np.int64) - gen.pairs[:, 0]).max() <= 20
    for i in range(len(rs)):
        compare_expectations(fe.transitions[i], fe.likelihood[i], ge.transitions[i], ge.likelihood[i])
    pm = port_model_from("ACGT", 6, trans, table)
    for i in (0, 17, 47):
        ora, _ = oracle_align(pm, reads[i], port_params(p))
        assert compare_pairs(fast.read_pairs(i), ora, p.threshold) <= 20


def test_pipelined_call_equals_single_batch():
    """sa_gpu_align_batch cuts large batches into chunks that overlap on the device; the result must be byte-identical
    to the un-chunked call for any chunk count (including more chunks than reads and an empty read in the middle)."""
    reads, table = synth.make_batch(23, seed=31, mean_len=700, min_len=200, max_len=1500, trim=0)
    model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=200)
    rs = [sa_read(r) for r in reads]
    rs.insert(11, sa.Read("", np.zeros((0, 4)), np.zeros((0, 2))))
    outs = []
    for nch in ("1", "3", "7", "64"):
        os.environ["SA_CHUNKS"] = nch
        try:
            r = sa.align_batch(model, rs, p)
        finally:
            os.environ.pop("SA_CHUNKS", None)
        outs.append((r.pair_offsets.copy(), r.pairs.copy(), r.band_cells, r.events))
    for o in outs[1:]:
        assert np.array_equal(o[0], outs[0][0]) and np.array_equal(o[1], outs[0][1]) and o[2:] == outs[0][2:]
    assert outs[0][0][12] == outs[0][0][11]     # the empty read has no pairs


def test_uneven_chunks_and_shutdown_keep_results():
    """Relative chunk sizes (SA_CHUNK_FRACS, the experiment knob of the pipelined call) do not change the result, and
    sa_gpu_shutdown gives the cached pinned / pool memory back without breaking a result the caller still holds or the
    next call."""
    reads, table = synth.make_batch(17, seed=77, mean_len=600, min_len=200, max_len=1200, trim=0)
    model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=200)
    rs = [sa_read(r) for r in reads]
    base = sa.align_batch(model, rs, p)
    want = (base.pair_offsets.copy(), base.pairs.copy())
    for fr in ("5,1", "1,2,3", "1,1,1,1,1,1,1,1,1,1,1,1,1,1,1,1,1,1,1,1"):
        os.environ["SA_CHUNK_FRACS"] = fr
        try:
            r = sa.align_batch(model, rs, p)
        finally:
            os.environ.pop("SA_CHUNK_FRACS", None)
        assert np.array_equal(r.pair_offsets, want[0]) and np.array_equal(r.pairs, want[1]), fr
        del r
    sa.gpu_shutdown()
    assert np.array_equal(base.pairs, want[1])          # the held result's slab survives the shutdown
    again = sa.align_batch(model, rs, p)
    assert np.array_equal(again.pair_offsets, want[0]) and np.array_equal(again.pairs, want[1])
