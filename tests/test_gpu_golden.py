"""-m gpu: the CUDA path, called through the C-ABI, against the GOLDEN outputs of the compiled reference.

Same fixtures as tests/test_oracle_golden.py (reference tests/stateMachineTests.c:842-983 known answers and the R9.4
10 922-event read), so on the GPU box — where neither /root/reference nor the compiled reference exists — parity is
still anchored on the reference's own outputs, not only on the oracle port.

Bars (BASELINE.json north_star): (x, y, path-kmer) bit-exact and in the reference's list order; every posterior within
1e-4 (1000 units of floor(p * 1e7)); pair set identical except within tolerance of the threshold; likelihood within
1e-5 relative; transition expectations within 1e-6 relative.  The FP64 kernels are in practice within a few units.
"""
import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port

from helpers import compare_expectations, compare_pairs, golden, golden_models, params_from

pytestmark = pytest.mark.gpu

TIGHT = 20   # observed bound of the FP64 kernels, in floor(p * 1e7) units (contract: 1000)


def _read(g, target, tp, rl=True, rr=True, events=None):
    return sa.Read(target, g["events"] if events is None else events, g["anchors"], float(tp[0]), float(tp[1]), float(tp[2]),
                   rl, rr)


def _align(model, reads, par, ambig=None, force_general=False):
    import os
    p = sa.default_params(**par)
    if force_general:
        os.environ["SA_FORCE_GENERAL"] = "1"
    try:
        return sa.align_batch(model, reads, p, ambig)
    finally:
        os.environ.pop("SA_FORCE_GENERAL", None)


def test_zymo_scaled_1076():
    g = golden("zymo_scaled")
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_NOISE_SCALED, g["tparams"])
    par = params_from(g["params"])
    res = _align(model, [_read(g, str(g["target"]), [1.0, 0.0, 1.0])], par)
    got = res.read_pairs(0)
    assert len(got) == 1076
    assert compare_pairs(got, g["pairs"], par["threshold"]) <= TIGHT


def test_zymo_descaled_1076_1119_and_degenerate_7349():
    g = golden("zymo_descaled")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_NOISE_DESCALE, tp)
    par = params_from(g["params"])
    t = str(g["target"])
    res = _align(model, [_read(g, t, tp)], par)
    assert len(res.read_pairs(0)) == 1076 and res.band_cells == 140469
    assert compare_pairs(res.read_pairs(0), g["pairs"], 0.01) <= TIGHT
    res = _align(model, [_read(g, t, tp)], dict(par, threshold=0.005))
    assert len(res.read_pairs(0)) == 1119
    assert compare_pairs(res.read_pairs(0), g["pairs_thr005"], 0.005) <= TIGHT
    # the degenerate-nucleotide test (stateMachineTests.c:920-983): one batch, four references, ragged ends off
    dst = ["C", "E", "O", "L"]
    res = _align(model, [_read(g, t.replace("C", d), tp, False, False) for d in dst], par)
    for i, (d, want) in enumerate(zip(dst, (1076, 1076, 1076, 7349))):
        got = res.read_pairs(i)
        assert len(got) == want, (d, len(got))
        assert compare_pairs(got, g["pairs_C2" + d], 0.01) <= TIGHT
    assert res.path_cells > res.band_cells
    # expectations with the inverse-Gaussian-noise emission
    e = sa.expectations_batch(model, [_read(g, t, tp)], sa.default_params(**par))
    compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions"], float(g["exp_likelihood"]), pseudo=0.001)


@pytest.mark.parametrize("force_general", [False, True])
def test_zymo_production_emission_all_parameter_sets(force_general):
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    t = str(g["target"])
    for tag in ("cli", "lib", "tight", "split"):
        par = params_from(g["params_" + tag])
        res = _align(model, [_read(g, t, tp)], par, force_general=force_general)
        assert compare_pairs(res.read_pairs(0), g["pairs_" + tag], par["threshold"]) <= TIGHT, tag
        e = sa.expectations_batch(model, [_read(g, t, tp)], sa.default_params(**par))
        compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions_" + tag],
                             float(g["exp_likelihood_" + tag]), pseudo=0.001)


def test_zymo_variant_positions_multipath():
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params_lib"])
    t = str(g["target_variants"])
    res = _align(model, [_read(g, t, tp)], par)
    assert compare_pairs(res.read_pairs(0), g["pairs_variants"], 0.01) <= TIGHT
    assert res.path_cells > res.band_cells
    e = sa.expectations_batch(model, [_read(g, t, tp)], sa.default_params(**par))
    compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions_variants"],
                         float(g["exp_likelihood_variants"]), pseudo=0.001)


def test_zymo_hdp_1217_and_assignments():
    g = golden("zymo_hdp")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_HDP, hdp=True)
    par = params_from(g["params"])
    res = _align(model, [_read(g, str(g["target"]), tp)], par)
    assert len(res.read_pairs(0)) == 1217
    assert compare_pairs(res.read_pairs(0), g["pairs"], par["threshold"]) <= TIGHT
    res = _align(model, [_read(g, str(g["target_variants"]), tp)], dict(par, threshold=0.01))
    assert compare_pairs(res.read_pairs(0), g["pairs_variants"], 0.01) <= TIGHT
    e = sa.expectations_batch(model, [_read(g, str(g["target"]), tp)], sa.default_params(**dict(par, threshold=0.01)))
    compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions"], float(g["exp_likelihood"]), pseudo=0.001)
    # HDP assignments (kmer offset in the target, event index), reference order (continuousHmm.c:510-514)
    a0, a1 = e.assign_offsets[0], e.assign_offsets[1]
    assert np.array_equal(e.assign_kmer_pos[a0:a1], g["exp_assign_kmer"])
    assert np.array_equal(e.assign_event[a0:a1], g["exp_assign_event"])


@pytest.mark.parametrize("force_general", [False, True])
def test_r9p4_selfref_10922_events(force_general):
    g = golden("r9p4_selfref")
    _, model = golden_models("model_r94_acgt", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params"])
    rd = sa.Read(str(g["target"]), g["events"], g["anchors"], *[float(v) for v in g["scaling"]])
    res = _align(model, [rd], par, force_general=force_general)
    got = res.read_pairs(0)
    assert len(got) == len(g["pairs"]) and res.events == 10922
    assert compare_pairs(got, g["pairs"], 0.01) <= TIGHT
    if not force_general:
        e = sa.expectations_batch(model, [rd], sa.default_params(**par))
        compare_expectations(e.transitions[0], e.likelihood[0], g["exp_transitions"], float(g["exp_likelihood"]), pseudo=0.001)


def test_mixed_batch_keeps_read_order_and_sums():
    """One batch holding reads of both models' shapes is not possible (one model per batch); but reads of very different
    sizes, ragged flags and an EMPTY read must come back in input order with per-read offsets intact."""
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params_lib"])
    t = str(g["target"])
    empty = sa.Read("", np.zeros((0, 4)), np.zeros((0, 2)))
    short = sa.Read(t[:5], g["events"][:3], np.zeros((0, 2)), float(tp[0]), float(tp[1]), float(tp[2]))   # lX = 0
    reads = [_read(g, t, tp), empty, _read(g, str(g["target_variants"]), tp), short, _read(g, t, tp)]
    res = _align(model, reads, par)
    assert res.n_reads == 5
    assert compare_pairs(res.read_pairs(0), g["pairs_lib"], 0.01) <= TIGHT
    assert len(res.read_pairs(1)) == 0 and len(res.read_pairs(3)) == 0
    assert compare_pairs(res.read_pairs(2), g["pairs_variants"], 0.01) <= TIGHT
    assert np.array_equal(res.read_pairs(4), res.read_pairs(0))
    e = sa.expectations_batch(model, reads, sa.default_params(**par))
    tab = np.zeros(10)
    import ctypes as C
    sa.lib().sa_result_sum_expectations(e._ptr, tab.ctypes.data_as(sa.c_f64p))
    # the lX = 0 read has total probability -inf, so (like the reference's exp(-inf - -inf)) its M->Y / Y->Y entries are
    # NaN and its likelihood -inf: the oracle port returns exactly that, and sums propagate it
    assert np.allclose(tab[:9], e.transitions.sum(0), equal_nan=True) and tab[9] == e.likelihood.sum()
    assert np.isnan(e.transitions[3][[2, 8]]).all() and e.likelihood[3] == -np.inf and not np.isnan(e.transitions[[0, 2, 4]]).any()
    assert np.array_equal(e.transitions[0], e.transitions[4]) and e.likelihood[0] == e.likelihood[4]


def test_bad_inputs_are_refused_not_computed():
    g = golden("zymo_meanonly")
    _, model = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    t = str(g["target"])
    with pytest.raises(sa.SaError, match="alphabet"):     # the reference exits in kmer_to_word (nanopore_hdp.c:393-398)
        sa.align_batch(model, [sa.Read(t[:100] + "N" + t[101:], g["events"], g["anchors"])])
    bad = g["anchors"].copy()
    bad[5] = bad[4]
    with pytest.raises(sa.SaError, match="anchor"):       # asserted by band_construct (pairwiseAligner.c:226-231)
        sa.align_batch(model, [sa.Read(t, g["events"], bad)])
    with pytest.raises(sa.SaError, match="parameters"):   # odd expansion (pairwiseAligner.c:1460-1464)
        sa.align_batch(model, [sa.Read(t, g["events"], g["anchors"])], sa.default_params(expansion=21))
