"""CPU suite, part 2: the C-ABI library without a device — exports, host-side integer logic, model loaders, errors.

No compute call is made here (there is no GPU in the build container); the compute entry points are only checked to
FAIL LOUDLY without a device (the product has no CPU fallback).
"""
import os
import re

import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port

from helpers import ROOT_DIR, golden, golden_models, scale_model_tables, scale_noise_tables, write_model_file, write_nhdp_file

REF = "/root/reference"
has_ref = os.path.exists(REF + "/models/testModelR73_acegot_template.model")


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT_DIR, "include", "signalalign_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(sa_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 35
    L = sa.lib()
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert declared == set(sa.EXPORTS), declared ^ set(sa.EXPORTS)


def test_compute_fails_loudly_without_device():
    if sa.lib().sa_gpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(sa.SaError, match="no CUDA device"):
        sa.gpu_init(0)
    m = golden("model_r94_acgt")
    model = sa.Model.from_tables("ACGT", 6, m["trans"], m["match"])
    r = sa.Read("ACGTACGTACGT", np.zeros((5, 4)), np.zeros((0, 2)))
    with pytest.raises(sa.SaError, match="no CPU fallback"):
        sa.align_batch(model, [r])
    with pytest.raises(sa.SaError, match="no CPU fallback"):
        sa.expectations_batch(model, [r])


def test_default_params_match_reference_constructor():
    p = sa.default_params()   # pairwiseAlignmentBandingParameters_construct, impl/pairwiseAligner.c:2022-2037
    assert (p.threshold, p.minDiagsBetweenTraceBack, p.traceBackDiagonals, p.diagonalExpansion,
            p.splitMatrixBiggerThanThis) == (0.01, 1000, 40, 20, 3000 * 3000)


def test_host_band_and_split_points_against_golden():
    g = golden("unit")
    for i in range(12):
        lX, lY, e = [int(v) for v in g["band%d_in" % i]]
        an = g["band%d_anchors" % i]
        assert np.array_equal(sa.debug_band(an, lX, lY, e), g["band%d_out" % i])
        for rl in (0, 1):
            for rr in (0, 1):
                got = sa.debug_split_points(an, lX, lY, int(g["split%d_bigger" % i]), rl, rr)
                assert np.array_equal(got, g["split%d_%d%d" % (i, rl, rr)])
    # tests/signalPairwiseAlignerTest.c:434-456
    got = sa.debug_band(np.array([[1, 0], [2, 1], [3, 3]]), 6, 5, 2)
    assert got[:, 1:].tolist() == [[0, 0], [-1, 1], [-2, 2], [-1, 3], [-2, 4], [-1, 3], [-2, 4], [-3, 3], [-2, 2], [-1, 3],
                                  [0, 2], [1, 1]]
    # empty inputs: no anchors, empty sequences
    assert sa.debug_band(np.zeros((0, 2)), 0, 0, 20).tolist() == [[0, 0, 0]]
    assert sa.debug_split_points(np.zeros((0, 2)), 7, 9, 10 ** 7, 1, 1).tolist() == [[0, 0, 7, 9]]


def test_host_band_random_against_oracle():
    rng = np.random.default_rng(5)
    for _ in range(200):
        lX, lY = int(rng.integers(0, 300)), int(rng.integers(0, 300))
        n = int(rng.integers(0, min(lX, lY) // 2 + 1))
        an = np.stack([np.sort(rng.choice(lX, n, replace=False)), np.sort(rng.choice(lY, n, replace=False))], 1) if n else np.zeros((0, 2))
        e = int(rng.choice([0, 2, 4, 10, 20, 100]))
        assert np.array_equal(sa.debug_band(an, lX, lY, e), port.band(an, lX, lY, e))
        big = int(rng.choice([1, 30, 500, 10 ** 6]))
        rl, rr = int(rng.integers(0, 2)), int(rng.integers(0, 2))
        assert np.array_equal(sa.debug_split_points(an, lX, lY, big, rl, rr), port.split_points(an, lX, lY, big, rl, rr))


def test_kmer_paths_order_and_errors(tmp_path):
    m = golden("model_r73_acegot")
    model = sa.Model.from_tables("ACEGOT", 6, m["trans"], m["match"])
    # default table: L = CEO (create_ambig_bases, impl/pairwiseAligner.c:32-65)
    p = model.kmer_paths("ATGLAL")
    assert [model.kmer_string(int(i)) for i in p] == ["ATGCAC", "ATGCAE", "ATGCAO", "ATGEAC", "ATGEAE", "ATGEAO",
                                                       "ATGOAC", "ATGOAE", "ATGOAO"]
    p = model.kmer_paths("LLLLLL")
    assert len(p) == 729 and model.kmer_string(int(p[0])) == "CCCCCC" and model.kmer_string(int(p[-1])) == "OOOOOO"
    assert model.kmer_paths("ACEGOT").tolist() == [model.kmer_index("ACEGOT")]
    with pytest.raises(sa.SaError):
        model.kmer_paths("ACGTNA")      # N is neither in the alphabet nor an ambiguity code
    assert len(model.kmer_paths("ACGTRA")) == 2   # R -> A / G, both in ACEGOT
    with pytest.raises(sa.SaError):
        model.kmer_paths("ACGTQA")      # Q -> A / I, and I is outside ACEGOT (the reference exits in kmer_to_word)
    # ambiguity file (create_ambig_bases2, impl/pairwiseAligner.c:68-92): code<TAB>bases
    f = tmp_path / "amb.positions"
    f.write_text("X\tCE\nq\tAT\n")
    amb = sa.Ambig(str(f))
    assert [model.kmer_string(int(i)) for i in model.kmer_paths("AXqAAA", amb)] == ["ACAAAA", "ACTAAA", "AEAAAA", "AETAAA"]
    # a file-loaded table replaces the defaults entirely: L is no longer ambiguous -> not in the alphabet
    with pytest.raises(sa.SaError):
        model.kmer_paths("ATGLAA", amb)


def test_kmer_string_roundtrip():
    m = golden("model_r94_rna5")
    model = sa.Model.from_tables(str(m["alphabet"]), int(m["k"]), m["trans"], m["match"])
    assert model.K == 1024 and model.k == 5
    for idx in (0, 1, 17, 1023):
        assert model.kmer_index(model.kmer_string(idx)) == idx
    assert model.kmer_string(0) == "AAAAA" and model.kmer_string(1023) == "TTTTT"


def test_model_file_loader_roundtrip(tmp_path):
    m = golden("model_r94_acgt")
    path = str(tmp_path / "t.model")
    write_model_file(path, "ACGT", 6, m["trans"], m["match"])
    model = sa.Model.load(path)
    t, mt, gy = model.tables()
    assert np.array_equal(mt, m["match"])
    assert np.array_equal(gy, m["gap_y"])               # 1.75 x sd on load (impl/stateMachine.c:1528-1532)
    fin = np.isfinite(m["trans"])
    assert np.allclose(t[fin], m["trans"][fin], rtol=0, atol=1e-15) and t[7] == -np.inf
    # malformed files are refused with a message (the reference st_errAbort()s)
    bad = tmp_path / "bad.model"
    bad.write_text("3\t4\tACGT\t6\n0.1\t0.2\n1.0\n")
    with pytest.raises(sa.SaError, match="transitions"):
        sa.Model.load(str(bad))
    with pytest.raises(sa.SaError, match="cannot open"):
        sa.Model.load(str(tmp_path / "missing.model"))


def test_scale_model_and_noise_match_restatement():
    m = golden("model_r94_acgt")
    model = sa.Model.from_tables("ACGT", 6, m["trans"], m["match"])
    model.scale_model(1.04, -2.2, 1.1, 0.9, 1.2)
    _, mt, gy = model.tables()
    em, ey = scale_model_tables(m["match"], m["gap_y"], 1.04, -2.2, 1.1, 0.9, 1.2)
    # numpy's pow and libm's pow may differ in the last ulp of the noise-sd column
    assert np.allclose(mt, em, rtol=4e-16, atol=0) and np.allclose(gy, ey, rtol=4e-16, atol=0)
    assert np.array_equal(mt[:, [0, 1, 2, 4]], em[:, [0, 1, 2, 4]])
    model2 = sa.Model.from_tables("ACGT", 6, m["trans"], m["match"])
    model2.scale_noise(0.9, 1.2)
    _, mt, gy = model2.tables()
    em, ey = scale_noise_tables(m["match"], m["gap_y"], 0.9, 1.2)
    assert np.allclose(mt, em, rtol=4e-16, atol=0) and np.allclose(gy, ey, rtol=4e-16, atol=0)


def test_nhdp_loader_and_density(tmp_path):
    """A small 2-level HDP in the reference's text format: leaves 0..15 (2-mers over ACGT), middle DPs 16..19, base 20.
    Only some leaves carry data; unobserved leaves must resolve to their first observed ancestor (dir_proc_density's
    parent walk, impl/hdp.c:2588-2612); the spline follows impl/hdp_math_utils.c:471-495."""
    rng = np.random.default_rng(1)
    G = 40
    grid = np.linspace(50.0, 120.0, G)
    parents = [16 + i // 4 for i in range(16)] + [20, 20, 20, 20, -1]
    data_ids = [0, 0, 5, 7, 7, 7]           # observed leaves 0, 5, 7 -> observed middles 16, 17 -> base 20
    pp, sl = {}, {}
    for dp in range(21):
        mu = rng.uniform(70, 100)
        pp[dp] = np.exp(-0.5 * ((grid - mu) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
        sl[dp] = np.gradient(pp[dp], grid)
    table = np.zeros((16, 5))
    table[:, 0] = 80.0
    table[:, 1] = 2.0
    table[:, 2] = 1.0
    table[:, 3] = 0.5
    table[:, 4] = 4.0
    trans = golden("model_r94_acgt")["trans"]
    mp, hp = str(tmp_path / "m.model"), str(tmp_path / "m.nhdp")
    write_model_file(mp, "ACGT", 2, trans, table)
    write_nhdp_file(hp, "ACGT", 2, grid, pp, sl, parents, data_ids)
    model = sa.Model.load(mp, hp)
    g2, k2r, lpp, lsl, _ = model.hdp_tables()
    assert np.array_equal(g2, grid)   # linspace restated exactly (impl/hdp_math_utils.c:497-510)
    # the shim's route: tables from the StateMachine struct, then the .nhdp attached (sa_model_attach_nhdp)
    attached = sa.Model.from_tables("ACGT", 2, trans, table)
    attached.attach_nhdp(hp)
    for a, b in zip(attached.hdp_tables(), model.hdp_tables()):
        assert np.array_equal(a, b)
    with pytest.raises(RuntimeError):
        attached.attach_nhdp(str(tmp_path / "absent.nhdp"))
    assert lpp.shape == (6, G)        # observed DPs: 0, 5, 7, 16, 17, 20
    resolve = {0: 0, 5: 5, 7: 7, 1: 16, 4: 17, 6: 17, 9: 20, 15: 20}
    pm = port.PortModel("ACGT", 2, trans, table, kind=port.EMIT_HDP, grid=g2, kmer_to_row=k2r, post_pred=lpp, slopes=lsl)
    pm.scale, pm.shift, pm.var = 1.0, 0.0, 1.0
    for kmer, dp in resolve.items():
        assert np.array_equal(lpp[k2r[kmer]], pp[dp]) and np.array_equal(lsl[k2r[kmer]], sl[dp])
        for x in (40.0, 50.0, 63.21, 88.8, 119.99, 120.0, 131.0):
            d = model.hdp_density(kmer, x)
            # independent restatement of grid_spline_interp
            if x <= grid[0]:
                want = pp[dp][0] - sl[dp][0] * (grid[0] - x)
            elif x >= grid[-1]:
                want = pp[dp][-1] + sl[dp][-1] * (x - grid[-1])
            else:
                dx = grid[1] - grid[0]
                i = int((x - grid[0]) / dx)
                dy = pp[dp][i + 1] - pp[dp][i]
                a, b = sl[dp][i] * dx - dy, dy - sl[dp][i + 1] * dx
                tl = (x - grid[i]) / dx
                tr = 1.0 - tl
                want = tr * pp[dp][i] + tl * pp[dp][i + 1] + tl * tr * (a * tr + b * tl)
            want = max(want, 0.0)
            assert d == want, (kmer, x, d, want)
    # expected values: only kmers whose OWN DP is observed are replaced (impl/stateMachine.c:1275-1304)
    model.set_hdp_expected_values()
    _, mt, _ = model.tables()
    changed = np.nonzero(mt[:, 0] != 80.0)[0].tolist()
    assert changed == [0, 5, 7]
    # alphabet mismatch between .model and .nhdp is refused (impl/stateMachine.c:1277-1288)
    write_nhdp_file(hp, "ACGU", 2, grid, pp, sl, parents, data_ids)
    with pytest.raises(sa.SaError, match="alphabet"):
        sa.Model.load(mp, hp)


def test_golden_hdp_tables_density_probes():
    g = golden("zymo_hdp")
    _, model = golden_models("model_r73_acegot", port.EMIT_HDP, hdp=True)
    for i, km in enumerate(g["probe_kmers"]):
        ki = model.kmer_index(str(km))
        for x, d in zip(g["probe_x"], g["probe_density"][i]):
            assert model.hdp_density(ki, float(x)) == d
    # stateMachine3_setModelToHdpExpectedValues against the reference's result (tests/stateMachineTests.c:874-900)
    model.set_hdp_expected_values()
    _, mt, _ = model.tables()
    assert np.array_equal(np.nonzero((mt[:, :2] != golden("model_r73_acegot")["match"][:, :2]).any(1))[0], g["changed_rows"])
    assert np.allclose(mt[:, :2], g["expected_value_cols"], rtol=1e-13, atol=0)


@pytest.mark.skipif(not has_ref, reason="/root/reference not present")
@pytest.mark.parametrize("name,fix", [("testModelR73_acegot_template.model", "model_r73_acegot"),
                                      ("testModelR9.4_450bps.nucleotide.6mer.template.model", "model_r94_acgt"),
                                      ("testModelR9p4_5mer_acgt_RNA.model", "model_r94_rna5")])
def test_real_model_files_load_identically(name, fix):
    g = golden(fix)
    model = sa.Model.load(REF + "/models/" + name)
    t, mt, gy = model.tables()
    assert model.alphabet == str(g["alphabet"]) and model.k == int(g["k"])
    assert np.array_equal(t, g["trans"]) and np.array_equal(mt, g["match"]) and np.array_equal(gy, g["gap_y"])


@pytest.mark.skipif(not has_ref, reason="/root/reference not present")
def test_real_nhdp_loads_identically():
    h = golden("hdp_single_level_fixed")
    model = sa.Model.load(REF + "/models/testModelR73_acegot_template.model", REF + "/models/templateSingleLevelFixed.nhdp")
    grid, k2r, pp, sl, own = model.hdp_tables()
    assert pp.shape == (352, 100)   # 352 observed DPs, grid 0..100 x 100
    assert np.array_equal(grid, h["grid"]) and np.array_equal(k2r, h["kmer_to_row"]) and np.array_equal(own, h["own_row"])
    assert np.array_equal(pp, h["post_pred"]) and np.array_equal(sl, h["slopes"])


def test_drop_in_build_exports_the_reference_entry_points():
    """oracle/_ref/libsaref_gpu.so (built where /root/reference exists): the reference's two entry points are provided by
    signalalign_b200/shim/signalalign_shim.c on top of libsagpu, the reference's own definitions are kept under cpu_*."""
    import shutil
    import subprocess
    path = os.path.join(ROOT_DIR, "oracle", "_ref", "libsaref_gpu.so")
    if not os.path.exists(path) or shutil.which("nm") is None:
        pytest.skip("drop-in build not present")
    out = subprocess.run(["nm", "-D", path], capture_output=True, text=True, check=True).stdout
    defined = {l.split()[-1] for l in out.splitlines() if " T " in l}
    undefined = {l.split()[-1] for l in out.splitlines() if " U " in l}
    for s in ("getAlignedPairsUsingAnchors", "getExpectationsUsingAnchors", "cpu_getAlignedPairsUsingAnchors",
              "cpu_getExpectationsUsingAnchors"):
        assert s in defined, s
    for s in ("sa_gpu_align_batch", "sa_gpu_expectations_batch", "sa_model_from_tables", "sa_kmer_string"):
        assert s in undefined, s
