"""variant-call rows and their per-position marginals.

CPU: the oracle restatement (oracle/port.py: variant_rows, marginalize_variants) against golden vectors produced by the
reference's own writePosteriorProbsVC and MarginalizeVariants (tests/golden/make_variants_golden.py).
GPU: sa_gpu_marginalize_variants through the C-ABI against the same goldens and against the oracle on a larger case."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

G = np.load(os.path.join(HERE, "golden", "variants.npz"))
N_CASES = int(G["n_cases"])


def case(i):
    k, comp, fwd, ev_off, ref_off = (int(v) for v in G["c%d_meta" % i])
    return dict(target=str(G["c%d_target" % i]), pairs=G["c%d_pairs" % i], kmers=[str(s) for s in G["c%d_kmers" % i]], k=k, comp=comp,
                fwd=fwd, ev_off=ev_off, ref_off=ref_off, variants=str(G["c%d_variants" % i]), tsv=str(G["c%d_tsv" % i]),
                positions=G["c%d_positions" % i], probs=G["c%d_probs" % i], per_read=G["c%d_per_read" % i])


def tsv_rows(text):
    rows = []
    for line in text.strip().split("\n"):
        f = line.split("\t")
        rows.append((int(f[0]), int(f[1]), f[2], float(f[3])))
    return rows


@pytest.mark.parametrize("i", range(N_CASES))
def test_oracle_rows_equal_the_reference_writer(i):
    from oracle import port
    c = case(i)
    rows = port.variant_rows(c["pairs"], c["target"], c["k"], c["kmers"], bool(c["comp"]), bool(c["fwd"]), c["ev_off"], c["ref_off"])
    want = tsv_rows(c["tsv"])
    assert len(rows) == len(want) > 0
    for r, w in zip(rows, want):
        assert r[:3] == w[:3]
        assert port.tsv_round6(r[3]) == w[3]          # the TSV carries six decimals


@pytest.mark.parametrize("i", range(N_CASES))
def test_oracle_marginals_equal_the_reference_class(i):
    from oracle import port
    c = case(i)
    rows = port.variant_rows(c["pairs"], c["target"], c["k"], c["kmers"], bool(c["comp"]), bool(c["fwd"]), c["ev_off"], c["ref_off"])
    pos, probs, per_read = port.marginalize_variants(rows, c["variants"])
    assert np.array_equal(pos, c["positions"])
    # (the reference sums in set-iteration order, which Python does not fix: equal up to rounding, not bit for bit)
    assert np.allclose(probs, c["probs"], rtol=1e-13, atol=0)
    assert np.allclose(per_read, c["per_read"], rtol=1e-12, atol=0)
    assert np.allclose(probs.sum(1), 1.0, rtol=1e-12)


# ------------------------------------------------------------------------------------------------------------ GPU
def _model(alphabet, k):
    import signalalign_b200 as sa
    from signalalign_b200 import synth
    K = len(alphabet) ** k
    table = np.zeros((K, 5))
    table[:, 0] = np.linspace(60, 120, K)
    table[:, 1] = 1.5
    table[:, 2] = 1.0
    table[:, 3] = 1.0
    table[:, 4] = 1.0
    return sa.Model.from_tables(alphabet, k, synth.nanopore_transitions(), table)


def _kmer_index(kmer, alphabet):
    idx = 0
    for c in kmer:
        idx = idx * len(alphabet) + alphabet.index(c)
    return idx


def _gpu_reads(cases, alphabet):
    reads = []
    for c in cases:
        pr = np.array(c["pairs"], dtype=np.int64).copy()
        pr[:, 3] = [_kmer_index(s, alphabet) for s in c["kmers"]]
        reads.append((pr.astype(np.int32), c["target"], c["comp"], c["fwd"], c["ref_off"]))
    return reads


@pytest.mark.gpu
def test_gpu_marginals_equal_the_reference_goldens():
    import signalalign_b200 as sa
    sa.gpu_init(0)
    alphabet = "ACEGOT"
    model = _model(alphabet, 6)
    for i in range(N_CASES):
        c = case(i)
        vs, res = sa.marginalize_variants(model, _gpu_reads([c], alphabet), c["variants"])
        pos, probs, per_read = res[0]
        assert vs == "".join(sorted(c["variants"]))
        assert np.array_equal(pos, c["positions"])
        assert np.allclose(probs, c["probs"], rtol=1e-12, atol=0)
        assert np.allclose(per_read, c["per_read"], rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_gpu_marginals_batch_equals_oracle_and_flags_bad_input():
    """all four golden cases in ONE call (one CTA per read), with and without the TSV rounding; then the error paths"""
    import signalalign_b200 as sa
    from oracle import port
    sa.gpu_init(0)
    alphabet = "ACEGOT"
    model = _model(alphabet, 6)
    cases = [case(i) for i in range(N_CASES)]
    for c in cases:
        c["variants"] = "ACEGOT"
    for rounding in (True, False):
        vs, res = sa.marginalize_variants(model, _gpu_reads(cases, alphabet), "ACEGOT", round_like_tsv=rounding)
        for c, (pos, probs, per_read) in zip(cases, res):
            rows = port.variant_rows(c["pairs"], c["target"], c["k"], c["kmers"], bool(c["comp"]), bool(c["fwd"]), c["ev_off"], c["ref_off"])
            opos, oprobs, oper = port.marginalize_variants(rows, "ACEGOT", round_like_tsv=rounding)
            assert np.array_equal(pos, opos)
            assert np.allclose(probs, oprobs, rtol=1e-12, atol=1e-300)
            assert np.allclose(per_read, oper, rtol=1e-12, atol=1e-300)
    # a read without pairs and a target without X: empty results, no error
    vs, res = sa.marginalize_variants(model, [(np.zeros((0, 4), np.int32), "ACGTACGTAC", 0, 1, 0),
                                              (np.array([[5000000, 1, 3, 0]], np.int32), "ACGTACGTAC", 0, 1, 0)], "CE")
    assert all(len(r[0]) == 0 for r in res) and np.all(res[0][2] == 0)
    # a called base outside the variants: the reference raises, so does the library
    c = case(0)
    with pytest.raises(sa.SaError):
        sa.marginalize_variants(model, _gpu_reads([c], alphabet), "C")
