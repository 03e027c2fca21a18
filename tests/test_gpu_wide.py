"""-m gpu: the WIDE band kernel (sa_wide.cuh: shared-memory DP state, any band width) through the C-ABI.

Three angles: (1) forced onto every read (SA_WIDE_ONLY) it must reproduce the oracle and the golden outputs of the
compiled reference exactly like the register-resident kernel — all parameter sets, ragged flags, expectations, HDP
emission and assignments, split regions, slot overflow; (2) with guide anchors derived the way signalMachine derives
them (match runs trimmed by 14 on both sides, SURVEY.md §8d: sparse anchors, bands hundreds to thousands of cells wide)
the default dispatch must pick it — no read may fall to the general kernel; (3) the same at the benchmark's read size.
"""
import os

import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port
from signalalign_b200 import synth

from helpers import (compare_expectations, compare_pairs, golden, golden_models, oracle_align, oracle_expectations,
                     params_from, port_model_from, port_params, sa_read)

pytestmark = pytest.mark.gpu
TIGHT = 20


class wide_only:
    def __enter__(self):
        os.environ["SA_WIDE_ONLY"] = "1"

    def __exit__(self, *a):
        os.environ.pop("SA_WIDE_ONLY", None)


def _batch(model, reads, p, expectations=False):
    b = sa.Batch(model, reads, p, expectations=expectations)
    kernels, wide = b.kernel_items, b.wide_items
    b.run()
    res = b.fetch()
    b.close()
    return res, kernels, wide


def _check_alignment(reads, table, p, wide=True, max_dp=10):
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    pm = port_model_from("ACGT", 6, trans, table)
    if wide:
        with wide_only():
            res, kernels, _ = _batch(model, [sa_read(r) for r in reads], p)
        assert kernels[3] == len(reads), kernels
    else:
        res, kernels, _ = _batch(model, [sa_read(r) for r in reads], p)
    cells, worst = 0, 0
    for i, r in enumerate(reads):
        ora, stats = oracle_align(pm, r, port_params(p))
        cells += stats[0]
        worst = max(worst, compare_pairs(res.read_pairs(i), ora, p.threshold))
    assert res.band_cells == cells
    assert worst <= max_dp, worst
    return kernels


@pytest.fixture(scope="module")
def dense_batch():
    return synth.make_batch(6, seed=11, mean_len=500, min_len=200, max_len=900, trim=0)


@pytest.fixture(scope="module")
def gapped_batch():
    return synth.make_batch(6, seed=12, mean_len=500, min_len=200, max_len=900, trim=14)


@pytest.fixture(scope="module")
def sparse_batch():
    # the contract's anchor pattern at a size the oracle finishes in seconds
    return synth.make_batch(5, seed=21, mean_len=1500, min_len=800, max_len=2500, **synth.ANCHOR_PATTERNS["contract"])


@pytest.mark.parametrize("expansion,trace_back,min_diags", [(50, 50, 200), (100, 50, 300), (20, 40, 1000), (6, 2, 10)])
def test_wide_kernel_forced_dense_and_gapped(dense_batch, gapped_batch, expansion, trace_back, min_diags):
    p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
    for reads, table in (dense_batch, gapped_batch):
        _check_alignment(reads, table, p)


def test_wide_kernel_ragged_flags_thresholds_and_slot_overflow(dense_batch):
    reads, table = dense_batch
    for rl, rr, thr in [(False, False, 0.01), (True, False, 0.2), (False, True, 0.001), (True, True, 1e-7)]:
        rs = [dict(r, ragged_left=rl, ragged_right=rr) for r in reads[:3]]
        p = sa.default_params(threshold=thr, expansion=50, trace_back=50, min_diags=200)
        _check_alignment(rs, table, p)       # 1e-7: more pairs than the first slot size -> grown and re-run


def test_sparse_anchors_default_dispatch_is_the_wide_kernel(sparse_batch):
    reads, table = sparse_batch
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    kernels = _check_alignment(reads, table, p, wide=False)
    assert kernels[0] == 0 and kernels[3] > 0, kernels
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=300)
    kernels = _check_alignment(reads, table, p, wide=False)
    assert kernels[0] == 0, kernels


@pytest.mark.parametrize("expansion,trace_back,min_diags", [(50, 50, 200), (100, 50, 300), (6, 2, 10)])
def test_wide_kernel_expectations(dense_batch, sparse_batch, expansion, trace_back, min_diags):
    trans = synth.nanopore_transitions()
    for reads, table in (dense_batch, sparse_batch):
        model = sa.Model.from_tables("ACGT", 6, trans, table)
        pm = port_model_from("ACGT", 6, trans, table)
        p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
        rs = [dict(r, ragged_left=(i % 2 == 0), ragged_right=(i % 3 != 0)) for i, r in enumerate(reads[:4])]
        with wide_only():
            res, kernels, _ = _batch(model, [sa_read(r) for r in rs], p, expectations=True)
        assert kernels[3] == len(rs), kernels
        for i, r in enumerate(rs):
            t, lik = oracle_expectations(pm, r, port_params(p))[:2]
            compare_expectations(res.transitions[i], res.likelihood[i], t, lik)


def _synthetic_hdp(table, seed=3, G=120):
    rng = np.random.default_rng(seed)
    K = table.shape[0]
    grid = np.linspace(40.0, 150.0, G)
    rows = 64
    mu = rng.uniform(60, 125, rows)
    sd = rng.uniform(1.5, 4.0, rows)
    pp = np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / sd[:, None]) ** 2) / (sd[:, None] * np.sqrt(2 * np.pi))
    slopes = np.gradient(pp, grid, axis=1)
    k2r = rng.integers(0, rows, K).astype(np.int32)
    return grid, k2r, pp, slopes


def test_wide_kernel_hdp_alignment_and_assignments(gapped_batch):
    reads, table = gapped_batch
    trans = synth.nanopore_transitions()
    grid, k2r, pp, slopes = _synthetic_hdp(table)
    model = sa.Model.from_tables("ACGT", 6, trans, table, kind=sa.EMIT_HDP)
    model.set_hdp(grid, k2r, pp, slopes)
    pm = port_model_from("ACGT", 6, trans, table, kind=port.EMIT_HDP, grid=grid, kmer_to_row=k2r, post_pred=pp, slopes=slopes)
    for expansion, trace_back, min_diags in [(50, 50, 200), (20, 10, 60)]:
        p = sa.default_params(threshold=0.01, expansion=expansion, trace_back=trace_back, min_diags=min_diags)
        rs = reads[:4]
        with wide_only():
            res, kernels, _ = _batch(model, [sa_read(r) for r in rs], p)
            exp, kernels_e, _ = _batch(model, [sa_read(r) for r in rs], p, expectations=True)
        assert kernels[3] == len(rs) and kernels_e[3] == len(rs), (kernels, kernels_e)
        for i, r in enumerate(rs):
            ora, _ = oracle_align(pm, r, port_params(p))
            assert compare_pairs(res.read_pairs(i), ora, p.threshold) <= TIGHT
            t, lik, ak, ae = oracle_expectations(pm, r, port_params(p))
            compare_expectations(exp.transitions[i], exp.likelihood[i], t, lik)
            a0, a1 = exp.assign_offsets[i], exp.assign_offsets[i + 1]
            assert np.array_equal(exp.assign_kmer_pos[a0:a1], ak) and np.array_equal(exp.assign_event[a0:a1], ae)


def test_wide_kernel_golden_zymo_all_parameter_sets_and_hdp_1217():
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    t = str(g["target"])
    rd = lambda: sa.Read(t, g["events"], g["anchors"], float(tp[0]), float(tp[1]), float(tp[2]))
    with wide_only():
        for tag in ("cli", "lib", "tight", "split"):
            par = params_from(g["params_" + tag])
            res, kernels, _ = _batch(model, [rd()], sa.default_params(**par))
            assert kernels[0] == 0 and kernels[3] >= 1, (tag, kernels)
            assert compare_pairs(res.read_pairs(0), g["pairs_" + tag], par["threshold"]) <= TIGHT, tag
            e, kernels, _ = _batch(model, [rd()], sa.default_params(**par), expectations=True)
            assert kernels[0] == 0, (tag, kernels)
            compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions_" + tag],
                                 float(g["exp_likelihood_" + tag]), pseudo=0.001)
        h = golden("zymo_hdp")
        tp = h["tparams"]
        _, hmodel = golden_models("model_r73_acegot", port.EMIT_HDP, hdp=True)
        par = params_from(h["params"])
        hr = lambda: sa.Read(str(h["target"]), h["events"], h["anchors"], float(tp[0]), float(tp[1]), float(tp[2]))
        res, kernels, _ = _batch(hmodel, [hr()], sa.default_params(**par))
        assert kernels[3] == 1, kernels
        assert len(res.read_pairs(0)) == 1217
        assert compare_pairs(res.read_pairs(0), h["pairs"], par["threshold"]) <= TIGHT
        e, kernels, _ = _batch(hmodel, [hr()], sa.default_params(**dict(par, threshold=0.01)), expectations=True)
        assert kernels[3] == 1, kernels
        compare_expectations(e.transitions[0], e.likelihood[0], h["exp_transitions"], float(h["exp_likelihood"]), pseudo=0.001)
        assert np.array_equal(e.assign_kmer_pos, h["exp_assign_kmer"]) and np.array_equal(e.assign_event, h["exp_assign_event"])


def test_wide_kernel_golden_r9p4_10922_events():
    g = golden("r9p4_selfref")
    _, model = golden_models("model_r94_acgt", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params"])
    rd = sa.Read(str(g["target"]), g["events"], g["anchors"], *[float(v) for v in g["scaling"]])
    with wide_only():
        res, kernels, _ = _batch(model, [rd], sa.default_params(**par))
        assert kernels[3] == 1, kernels
        assert compare_pairs(res.read_pairs(0), g["pairs"], 0.01) <= TIGHT
        e, _, _ = _batch(model, [rd], sa.default_params(**par), expectations=True)
    compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions"], float(g["exp_likelihood"]), pseudo=0.001)


def test_wide_kernel_equals_register_kernel_byte_for_byte(dense_batch):
    """same arithmetic, same accumulation order: the two band kernels must agree on every pair, not only within tolerance"""
    reads, table = synth.make_batch(24, seed=31, mean_len=1200, min_len=600, max_len=2000, trim=0)
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    rs = [sa_read(r) for r in reads]
    a, ka, _ = _batch(model, rs, p)
    with wide_only():
        b, kb, _ = _batch(model, rs, p)
    assert ka[3] == 0 and kb[3] == len(reads), (ka, kb)
    assert np.array_equal(a.pair_offsets, b.pair_offsets) and np.array_equal(a.pairs, b.pairs)


def test_contract_anchor_pattern_at_benchmark_read_size():
    """~6 k-nt / ~10 k-event reads, anchors from match runs trimmed by 14, -x 100: bands up to
    ~2 500 cells wide.  Every read must run on the wide kernel (none on the general kernel) and match the oracle."""
    import bench
    reads, table = synth.make_batch(3, seed=20261104, **bench.GEN)
    p = sa.default_params(threshold=bench.THRESHOLD, expansion=bench.EXPANSION, trace_back=bench.TRACE_BACK, min_diags=bench.MIN_DIAGS)
    kernels = _check_alignment(reads, table, p, wide=False)
    assert kernels == (0, 0, 0, len(reads)), kernels
    # the CLI-default pattern (long match runs trimmed by 14): bands 100 - 250 wide
    reads, table = synth.make_batch(3, seed=20261104, **synth.ANCHOR_PATTERNS["long_runs_trim14"])
    kernels = _check_alignment(reads, table, p, wide=False)
    assert kernels[0] == 0, kernels
