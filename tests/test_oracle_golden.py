"""CPU suite, part 1: the oracle port (oracle/sa_oracle.c) pinned against the golden fixtures.

The fixtures in tests/golden/ were produced by the UNMODIFIED reference compiled from /root/reference
(tests/golden/make_golden.py); they include the reference's own known-answer counts (1076 / 1119 / 1217 / 7349,
tests/stateMachineTests.c:842-983).  The oracle must reproduce them bit for bit: same pairs, same order, same
floor(p * 1e7), same path kmers; expectations to 1e-12 relative.
"""
import math

import numpy as np
import pytest

from oracle import port

from helpers import golden, golden_models, params_from


def _kmer_idx(kmer, alphabet):
    v = 0
    for ch in kmer:
        v = v * len(alphabet) + alphabet.index(ch)
    return v


def _align(pm, g, target, par, tp, rl=True, rr=True):
    pm.scale, pm.shift, pm.var = float(tp[0]), float(tp[1]), float(tp[2])
    pairs, stats = port.align(pm, target, g["events"], g["anchors"], port.params(**par), ragged_left=rl, ragged_right=rr)
    return pairs, stats


def test_logadd_bit_exact():
    g = golden("unit")
    got = np.array([port.log_add(float(a), float(b)) for a, b in zip(g["logadd_x"], g["logadd_y"])])
    assert np.array_equal(got.view(np.int64), g["logadd_out"].view(np.int64))
    # the reference's own tolerance test (tests/signalPairwiseAlignerTest.c:115-125): within 1e-3 of log(e^x + e^y)
    fin = np.isfinite(g["logadd_x"]) & np.isfinite(g["logadd_y"])
    exact = np.logaddexp(g["logadd_x"][fin], g["logadd_y"][fin])
    assert np.abs(got[fin] - exact).max() < 1e-3


def test_kmer_id_known_answers():
    g = golden("unit")
    alpha = "ACEGOT"
    got = np.array([port.kmer_id(str(k), alpha) for k in g["emit_kmers"]])
    assert np.array_equal(got, g["kmer_ids"])
    # tests/nanoporeHdpTests.c:102-108 style known answers
    assert port.kmer_id("AAAAAA", "ACGT") == 0 and port.kmer_id("TTTTTT", "ACGT") == 4095
    assert port.kmer_id("AAAAAC", "ACGT") == 1 and port.kmer_id("CAAAAA", "ACGT") == 1024
    assert port.kmer_id("AAXAAA", "ACGT") == -1


@pytest.mark.parametrize("tag,kind", [("meanonly", port.EMIT_GAUSS_MEANONLY_DESCALE), ("scaled", port.EMIT_GAUSS_NOISE_SCALED),
                                      ("descale", port.EMIT_GAUSS_NOISE_DESCALE)])
def test_gaussian_emissions_bit_exact(tag, kind):
    g = golden("unit")
    m = golden("model_r73_acegot")
    pm = port.PortModel(str(m["alphabet"]), int(m["k"]), m["trans"], m["match"], gap_y=m["gap_y"], gap_x=m["gap_x"], kind=kind)
    pm.scale, pm.shift, pm.var = [float(v) for v in g["emit_scaling"]]
    want = g["emit_" + tag]
    for i, km in enumerate(g["emit_kmers"]):
        ki = _kmer_idx(str(km), pm.alphabet)
        for col, match in ((0, True), (1, False)):
            got = port.emission(pm, ki, float(g["emit_mean"][i]), float(g["emit_noise"][i]), match)
            assert got == want[i, col] or (np.isnan(got) and np.isnan(want[i, col])), (tag, km, match, got, want[i, col])


def test_model_file_quirks():
    """tests/stateMachineTests.c:233-291: gap-Y sd = 1.75 x match sd, other columns equal; GAP_SWITCH_TO_X stays -inf"""
    m = golden("model_r94_acgt")
    assert np.array_equal(m["gap_y"][:, 1], m["match"][:, 1] * 1.75)
    assert np.array_equal(m["gap_y"][:, [0, 2, 3, 4]], m["match"][:, [0, 2, 3, 4]])
    assert m["trans"][7] == -np.inf and np.all(m["gap_x"] == np.log(0.1))


def test_bands_and_split_points():
    g = golden("unit")
    for i in range(12):
        lX, lY, e = [int(v) for v in g["band%d_in" % i]]
        an = g["band%d_anchors" % i]
        assert np.array_equal(port.band(an, lX, lY, e), g["band%d_out" % i])
        for rl in (0, 1):
            for rr in (0, 1):
                got = port.split_points(an, lX, lY, int(g["split%d_bigger" % i]), rl, rr)
                assert np.array_equal(got, g["split%d_%d%d" % (i, rl, rr)])
    # the reference's exact-diagonal test (tests/signalPairwiseAlignerTest.c:434-456)
    want = [(0, 0, 0), (1, -1, 1), (2, -2, 2), (3, -1, 3), (4, -2, 4), (5, -1, 3), (6, -2, 4), (7, -3, 3), (8, -2, 2),
            (9, -1, 3), (10, 0, 2), (11, 1, 1)]
    assert port.band(np.array([[1, 0], [2, 1], [3, 3]]), 6, 5, 2).tolist() == [list(w) for w in want]


def test_hdcell_path_order():
    """tests/signalPairwiseAlignerTest.c:543-568: ATGXAX with X = CEO -> 9 paths ATGCAC .. ATGOAO; XXXXXX -> 729"""
    m = golden("model_r73_acegot")
    pm = port.PortModel(str(m["alphabet"]), 6, m["trans"], m["match"])
    mc = pm.cstruct()
    amb = port.make_ambig()
    amb.opts[ord("X")].value = b"CEO"
    import ctypes as C
    out = np.zeros(1024, np.int32)
    n = port.lib().sao_paths(b"ATGXAX", C.byref(mc), C.byref(amb), out.ctypes.data_as(port.c_i32p), 1024)
    assert n == 9 and pm.kmer_string(int(out[0])) == "ATGCAC" and pm.kmer_string(int(out[8])) == "ATGOAO"
    n = port.lib().sao_paths(b"XXXXXX", C.byref(mc), C.byref(amb), out.ctypes.data_as(port.c_i32p), 1024)
    assert n == 729 and pm.kmer_string(int(out[0])) == "CCCCCC" and pm.kmer_string(int(out[728])) == "OOOOOO"


def test_zymo_scaled_1076_and_unbanded():
    g = golden("zymo_scaled")
    pm, _ = golden_models("model_r73_acegot", port.EMIT_GAUSS_NOISE_SCALED, g["tparams"])
    pairs, _ = _align(pm, g, str(g["target"]), params_from(g["params"]), [1.0, 0.0, 1.0])
    assert len(pairs) == 1076 and np.array_equal(pairs, g["pairs"])
    pm.scale, pm.shift, pm.var = 1.0, 0.0, 1.0
    nb = port.align_noband(pm, str(g["target"]), g["events"], port.params(**params_from(g["params"])))
    assert len(nb) == 1076 and np.array_equal(nb[:, :3], g["pairs_noband"])


def test_zymo_descaled_1076_1119_and_degenerate_7349():
    g = golden("zymo_descaled")
    tp = g["tparams"]
    pm, _ = golden_models("model_r73_acegot", port.EMIT_GAUSS_NOISE_DESCALE, tp)
    par = params_from(g["params"])
    pairs, stats = _align(pm, g, str(g["target"]), par, tp)
    assert len(pairs) == 1076 and np.array_equal(pairs, g["pairs"])
    assert int(stats[0]) == 140469   # band cells of the probe read
    p5, _ = _align(pm, g, str(g["target"]), dict(par, threshold=0.005), tp)
    assert len(p5) == 1119 and np.array_equal(p5, g["pairs_thr005"])
    for dst, want in (("C", 1076), ("E", 1076), ("O", 1076), ("L", 7349)):
        q, _ = _align(pm, g, str(g["target"]).replace("C", dst), par, tp, rl=False, rr=False)
        assert len(q) == want and np.array_equal(q, g["pairs_C2" + dst]), dst
    pm.scale, pm.shift, pm.var = float(tp[0]), float(tp[1]), float(tp[2])
    tr, lik, _, _ = port.expectations(pm, str(g["target"]), g["events"], g["anchors"], port.params(**par))
    assert np.allclose(tr, g["exp_transitions"], rtol=1e-12, atol=0) and abs(lik / float(g["exp_likelihood"]) - 1) < 1e-12


@pytest.mark.parametrize("tag", ["cli", "lib", "tight", "split"])
def test_zymo_production_emission(tag):
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    pm, _ = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params_" + tag])
    pairs, _ = _align(pm, g, str(g["target"]), par, tp)
    assert np.array_equal(pairs, g["pairs_" + tag])
    tr, lik, _, _ = port.expectations(pm, str(g["target"]), g["events"], g["anchors"], port.params(**par))
    assert np.allclose(tr, g["exp_transitions_" + tag], rtol=1e-12, atol=0)
    assert abs(lik / float(g["exp_likelihood_" + tag]) - 1) < 1e-12


def test_zymo_variant_positions():
    g = golden("zymo_meanonly")
    tp = g["tparams"]
    pm, _ = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params_lib"])
    pairs, stats = _align(pm, g, str(g["target_variants"]), par, tp)
    assert np.array_equal(pairs, g["pairs_variants"])
    assert stats[1] > stats[0]   # path-cells exceed band cells: multi-path HDCells were exercised
    tr, lik, _, _ = port.expectations(pm, str(g["target_variants"]), g["events"], g["anchors"], port.params(**par))
    assert np.allclose(tr, g["exp_transitions_variants"], rtol=1e-12, atol=0)
    assert abs(lik / float(g["exp_likelihood_variants"]) - 1) < 1e-12


def test_zymo_hdp_1217_density_and_assignments():
    g = golden("zymo_hdp")
    tp = g["tparams"]
    pm, _ = golden_models("model_r73_acegot", port.EMIT_HDP, hdp=True)
    par = params_from(g["params"])
    pairs, _ = _align(pm, g, str(g["target"]), par, tp)
    assert len(pairs) == 1217 and np.array_equal(pairs, g["pairs"])
    pv, _ = _align(pm, g, str(g["target_variants"]), dict(par, threshold=0.01), tp)
    assert np.array_equal(pv, g["pairs_variants"])
    tr, lik, ak, ae = port.expectations(pm, str(g["target"]), g["events"], g["anchors"],
                                        port.params(**dict(par, threshold=0.01)))
    assert np.allclose(tr, g["exp_transitions"], rtol=1e-12, atol=0) and abs(lik / float(g["exp_likelihood"]) - 1) < 1e-12
    assert np.array_equal(ak, g["exp_assign_kmer"]) and np.array_equal(ae, g["exp_assign_event"])
    # density probes through the emission: log(1/var * density) at var = 1, scale = 1, shift = 0 -> log(density)
    pm.scale, pm.shift, pm.var = 1.0, 0.0, 1.0
    for i, km in enumerate(g["probe_kmers"]):
        ki = _kmer_idx(str(km), pm.alphabet)
        for x, d in zip(g["probe_x"], g["probe_density"][i]):
            got = port.emission(pm, ki, float(x), 1.0, True)
            want = math.log(d) if d > 0 else -math.inf   # libm log, as the reference calls it
            assert got == want or abs(got - want) < 1e-9, (km, x, got, want)  # de-scaling moves x by an ulp


def test_r9p4_selfref_10922_events():
    g = golden("r9p4_selfref")
    pm, _ = golden_models("model_r94_acgt", port.EMIT_GAUSS_MEANONLY_DESCALE)
    par = params_from(g["params"])
    pairs, stats = _align(pm, g, str(g["target"]), par, g["scaling"])
    assert g["events"].shape[0] == 10922 and np.array_equal(pairs, g["pairs"])
    tr, lik, _, _ = port.expectations(pm, str(g["target"]), g["events"], g["anchors"], port.params(**par))
    assert np.allclose(tr, g["exp_transitions"], rtol=1e-12, atol=0) and abs(lik / float(g["exp_likelihood"]) - 1) < 1e-12
