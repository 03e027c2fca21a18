"""adaptive banded event -> kmer alignment (reference impl/eventAligner.c:899-1235) and the method-of-moments
scalings (:792-842).

CPU: the numpy restatement (oracle/port.py: event_align, mom_scalings) against golden vectors produced by the reference's
own functions (tests/golden/make_event_align_golden.py).  GPU: sa_gpu_event_align_batch through the C-ABI against the
same goldens (bit-exact pairs) and against the oracle on fresh reads."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

G = np.load(os.path.join(HERE, "golden", "event_align.npz"))
N_CASES = int(G["n_cases"])
ALPHABET, K = str(G["alphabet"]), int(G["k"])
MATCH = np.asarray(G["match"], np.float64).reshape(-1, 5)


def kmer_ids(seq, rna):
    A = len(ALPHABET)
    out = []
    for i in range(len(seq) - K + 1):
        kmer = seq[i:i + K].replace("U", "T")
        if rna:
            kmer = kmer[::-1]                       # build_kmer_list, eventAligner.c:774-790
        idx = 0
        for c in kmer:
            idx = idx * A + ALPHABET.index(c)
        out.append(idx)
    return np.array(out, np.int64)


def case(i):
    rna, mom = (int(v) for v in G["c%d_flags" % i])
    return dict(seq=str(G["c%d_sequence" % i]), mean=G["c%d_mean" % i], rna=rna, mom=mom, scalings=G["c%d_scalings" % i],
                pairs=G["c%d_pairs" % i])


@pytest.mark.parametrize("i", range(N_CASES))
def test_oracle_equals_the_reference_aligner(i):
    from oracle import port
    c = case(i)
    kid = kmer_ids(c["seq"], c["rna"])
    if c["mom"]:
        scale, shift, var = port.mom_scalings(MATCH, kid, c["mean"])
        assert (scale, shift, var) == tuple(c["scalings"])          # bit for bit
    else:
        scale, shift, var = c["scalings"]
    pairs, qc = port.event_align(MATCH, kid, c["mean"], scale, shift, var)
    assert np.array_equal(pairs, c["pairs"])
    assert qc["failed"] == (len(c["pairs"]) == 0)


# ------------------------------------------------------------------------------------------------------------ GPU
def _model():
    import signalalign_b200 as sa
    from signalalign_b200 import synth
    return sa.Model.from_tables(ALPHABET, K, synth.nanopore_transitions(), MATCH)


@pytest.mark.gpu
def test_gpu_event_alignment_equals_the_reference_goldens():
    """all golden reads in ONE call: aligned pairs and the method-of-moments scalings bit for bit, QC verdicts equal"""
    import signalalign_b200 as sa
    sa.gpu_init(0)
    cases = [case(i) for i in range(N_CASES)]
    res = sa.event_align_batch(_model(), [dict(sequence=c["seq"], event_mean=c["mean"], rna=c["rna"],
                                               scalings=None if c["mom"] else tuple(c["scalings"])) for c in cases])
    for c, r in zip(cases, res):
        assert (r["scale"], r["shift"], r["var"]) == tuple(c["scalings"])
        assert np.array_equal(r["pairs"], c["pairs"])
        assert r["failed"] == (len(c["pairs"]) == 0)


@pytest.mark.gpu
def test_gpu_event_alignment_equals_oracle_on_fresh_reads_and_flags_bad_input():
    import signalalign_b200 as sa
    from signalalign_b200 import synth
    from oracle import port
    sa.gpu_init(0)
    rng = np.random.default_rng(77)
    ref_codes = rng.integers(0, len(ALPHABET), 80000).astype(np.int64)
    reads = []
    for _ in range(6):
        r = synth.make_read(rng, ref_codes, MATCH, ALPHABET, K, mean_len=3000, sigma_len=0.3, min_len=300, max_len=8000,
                            n_probs=(0.08, 0.62, 0.22, 0.08))
        reads.append(dict(sequence=r["target"], event_mean=r["events"][:, 0].astype(np.float32), rna=False, scalings=None))
    reads.append(dict(sequence="ACGTACG", event_mean=np.array([80.0, 90.0, 85.0], np.float32), rna=False, scalings=(1.0, 0.0, 1.0)))
    res = sa.event_align_batch(_model(), reads)
    for rd, r in zip(reads, res):
        kid = kmer_ids(rd["sequence"], False)
        sc = port.mom_scalings(MATCH, kid, rd["event_mean"]) if rd["scalings"] is None else rd["scalings"]
        assert (r["scale"], r["shift"], r["var"]) == tuple(sc)
        pairs, qc = port.event_align(MATCH, kid, rd["event_mean"], *sc)
        assert np.array_equal(r["pairs"], pairs)
        assert r["failed"] == qc["failed"] and r["max_gap"] == qc["max_gap"] and r["n_aligned_before_qc"] == qc["n_aligned"]
        assert r["avg_log_emission"] == qc["avg_log_emission"]
    with pytest.raises(sa.SaError):
        sa.event_align_batch(_model(), [dict(sequence="ACGTNACGTAC", event_mean=np.ones(4, np.float32))])
    with pytest.raises(sa.SaError):
        sa.event_align_batch(_model(), [dict(sequence="ACG", event_mean=np.ones(4, np.float32))])
