#!/usr/bin/env python
"""Generates the golden fixtures in this directory from the UNMODIFIED reference compiled into oracle/_ref/.

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden.py

Every fixture is a compressed .npz holding the exact inputs of one reference call (plain arrays) and the outputs the
reference produced for them, so the parity tests can run where /root/reference does not exist (the GPU box).  The
cases follow the reference's own known-answer tests:

  zymo_*          tests/stateMachineTests.c:842-983 (1076 / 1076 / 1076 un-banded / 1217 HDP / 1076, 1076, 7349 degenerate)
  r9p4_selfref    the 10 922-event R9.4 1-D fixture against its own base-called sequence (E. coli reference blob is
                  missing from the snapshot), production emission (signalMachine.c:347), CLI parameters -x 50 -m 14
  expect_*        getExpectationsUsingAnchors (tests/stateMachineTests.c:1003-1083 shape): transitions + likelihood
  unit            logAdd, Gaussian / inverse-Gaussian emissions, kmer_id, band_construct, getSplitPoints, HDCell paths
  models          the numeric tables of the model files the cases use (so the cases are self-contained)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

REF = "/root/reference"
M_R73 = REF + "/models/testModelR73_acegot_template.model"
M_R94 = REF + "/models/testModelR9.4_450bps.nucleotide.6mer.template.model"
M_RNA = REF + "/models/testModelR9p4_5mer_acgt_RNA.model"
NHDP = REF + "/models/templateSingleLevelFixed.nhdp"
ZYMO_READ = REF + "/tests/test_npReads/ZymoC_ch_1_file1.npRead"
ZYMO_REF = REF + "/tests/test_npReads/ZymoRef.txt"
R94_READ = REF + "/tests/test_npReads/r9p4_oneD.npRead"


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **kw)
    print("%-28s %8.1f KB" % (name + ".npz", os.path.getsize(path) / 1024.0))


def kmer_idx(kmers, alphabet):
    a = len(alphabet)
    lut = {c: i for i, c in enumerate(alphabet)}
    out = np.zeros(len(kmers), np.int64)
    for i, km in enumerate(kmers):
        v = 0
        for ch in km:
            v = v * a + lut[ch]
        out[i] = v
    return out


def pairs4(pairs, kmers, alphabet):
    return np.concatenate([pairs, kmer_idx(kmers, alphabet)[:, None]], axis=1)


def model_arrays(sm):
    m, y, x = sm.tables()
    return dict(alphabet=np.array(sm.alphabet), k=np.int64(sm.k), trans=sm.transitions.copy(), match=m, gap_y=y,
                gap_x=x, scaling=sm.scaling())


def main():
    rng = np.random.default_rng(20261104)

    # ------------------------------------------------------------------ model tables
    raw73 = ref.RefStateMachine(M_R73, kind=0)
    save("model_r73_acegot", **model_arrays(raw73))
    raw94 = ref.RefStateMachine(M_R94, kind=0)
    save("model_r94_acgt", **model_arrays(raw94))
    rawrna = ref.RefStateMachine(M_RNA, kind=0)
    save("model_r94_rna5", **model_arrays(rawrna))

    # ------------------------------------------------------------------ Zymo read (R7.3, ACEGOT)
    zref = open(ZYMO_REF).read().strip()
    rd = ref.RefNpRead(ZYMO_READ, "ACEGOT")
    tp = dict(rd.template_params)
    events = rd.template_events()
    anchors = ref.lastz_anchors(zref, rd.twoD_read, rd.template_event_map())
    common = dict(target=np.array(zref), events=events, anchors=anchors,
                  tparams=np.array([tp[k] for k in ("scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift")]))
    dflt = dict(threshold=0.01, min_diags=1000, trace_back=40, expansion=20, split_bigger=3000 * 3000)

    # (1) scaled model, emission incl. inverse-Gaussian noise (loadScaledStateMachine3)
    sm = ref.RefStateMachine(M_R73, kind=1)
    sm.scale_model(tp["scale"], tp["shift"], tp["var"], tp["scale_sd"], tp["var_sd"])
    p, km, _ = sm.align(zref, events, anchors, **dflt)
    assert len(p) == 1076, len(p)
    nb = sm.align_noband(zref, events, threshold=0.01)
    assert len(nb) == 1076, len(nb)
    save("zymo_scaled", kind=np.int64(2), pairs=pairs4(p, km, sm.alphabet), pairs_noband=nb, params=np.array(list(dflt.values())),
         ragged=np.array([1, 1]), **common)

    # (2) de-scaled emission with scaleNoise (loadDescaledStateMachine3)
    sm = ref.RefStateMachine(M_R73, kind=2)
    sm.set_scaling(tp["scale"], tp["shift"], tp["var"])
    sm.scale_noise(tp["scale_sd"], tp["var_sd"])
    p, km, _ = sm.align(zref, events, anchors, **dflt)
    assert len(p) == 1076, len(p)
    p005, km005, _ = sm.align(zref, events, anchors, **dict(dflt, threshold=0.005))
    assert len(p005) == 1119, len(p005)
    deg = {}
    for src, dst, want in (("C", "C", 1076), ("C", "E", 1076), ("C", "O", 1076), ("C", "L", 7349)):
        t2 = zref.replace(src, dst)
        q, kq, _ = sm.align(t2, events, anchors, ragged_left=False, ragged_right=False, **dflt)
        assert len(q) == want, (dst, len(q))
        deg["pairs_C2" + dst] = pairs4(q, kq, sm.alphabet)
    # expectations with this machine (getExpectationsUsingAnchors, pseudocounts 0.001)
    tr, lik, _, _, _ = sm.expectations(zref, events, anchors, **dflt)
    save("zymo_descaled", kind=np.int64(3), pairs=pairs4(p, km, sm.alphabet), pairs_thr005=pairs4(p005, km005, sm.alphabet),
         params=np.array(list(dflt.values())), ragged=np.array([1, 1]), exp_transitions=tr, exp_likelihood=np.float64(lik),
         **deg, **common)

    # (3) production emission (MeanOnly de-scaling, buildStateMachine2) — alignment, expectations, ambiguity, split
    sm = ref.RefStateMachine(M_R73, kind=0)
    sm.set_scaling(tp["scale"], tp["shift"], tp["var"])
    out = {}
    for tag, par in (("cli", dict(threshold=0.01, min_diags=1000, trace_back=50, expansion=50, split_bigger=3000 * 3000)),
                     ("lib", dflt),
                     ("tight", dict(threshold=0.05, min_diags=100, trace_back=10, expansion=10, split_bigger=3000 * 3000)),
                     ("split", dict(threshold=0.01, min_diags=60, trace_back=20, expansion=20, split_bigger=150))):
        p, km, _ = sm.align(zref, events, anchors, **par)
        out["pairs_" + tag] = pairs4(p, km, sm.alphabet)
        out["params_" + tag] = np.array(list(par.values()))
        tr, lik, _, _, _ = sm.expectations(zref, events, anchors, **par)
        out["exp_transitions_" + tag] = tr
        out["exp_likelihood_" + tag] = np.float64(lik)
    # cytosine variants at a handful of positions (variantCaller shape): X = ACGT..., here P = C/E and L = C/E/O
    cpos = [i for i, ch in enumerate(zref) if ch == "C"]
    pick = rng.choice(cpos, 12, replace=False)
    t2 = list(zref)
    for j, i in enumerate(sorted(pick)):
        t2[i] = "P" if j % 2 else "L"
    t2 = "".join(t2)
    p, km, _ = sm.align(t2, events, anchors, **dflt)
    out["target_variants"] = np.array(t2)
    out["pairs_variants"] = pairs4(p, km, sm.alphabet)
    tr, lik, _, _, _ = sm.expectations(t2, events, anchors, **dflt)
    out["exp_transitions_variants"] = tr
    out["exp_likelihood_variants"] = np.float64(lik)
    save("zymo_meanonly", kind=np.int64(0), **out, **common)

    # (4) HDP emissions on the de-scaled read (test_sm3Hdp_getAlignedPairsWithBanding: 1217 pairs at 0.1)
    rd2 = ref.RefNpRead(ZYMO_READ, "ACEGOT")
    rd2.descale()
    tp2 = dict(rd2.template_params)
    ev2 = rd2.template_events()
    smh = ref.RefStateMachine(M_R73, kind=0, nhdp_file=NHDP)
    smh.set_scaling(tp2["scale"], tp2["shift"], tp2["var"])
    p, km, _ = smh.align(zref, ev2, anchors, **dict(dflt, threshold=0.1))
    assert len(p) == 1217, len(p)
    p01, km01, _ = smh.align(t2, ev2, anchors, **dflt)
    tr, lik, ak, ae, _ = smh.expectations(zref, ev2, anchors, **dflt)
    # density probes (nanoporeHdpTests.c:464-478 probes x = 65 for five kmers; plus a sweep incl. both extrapolation sides)
    probes = ["AAAAAA", "GCACCA", "CCTTAG", "ACTTCA", "GGAATC", "CEOCEO", "TTTTTT", "OOOOOO", "ACEGOT"]
    xs = np.concatenate([[65.0], np.linspace(-5.0, 105.0, 45)])
    dens = np.array([[smh.hdp_density(k, float(x)) for x in xs] for k in probes])
    m, y, x = smh.tables()
    smh.hdp_expected_values()
    m2, _, _ = smh.tables()
    save("zymo_hdp", kind=np.int64(1), target=np.array(zref), events=ev2, anchors=anchors,
         tparams=np.array([tp2[k] for k in ("scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift")]),
         params=np.array(list(dict(dflt, threshold=0.1).values())), pairs=pairs4(p, km, smh.alphabet),
         target_variants=np.array(t2), pairs_variants=pairs4(p01, km01, smh.alphabet),
         exp_transitions=tr, exp_likelihood=np.float64(lik), exp_assign_kmer=ak, exp_assign_event=ae,
         probe_kmers=np.array(probes), probe_x=xs, probe_density=dens,
         expected_value_cols=m2[:, :2].copy(), changed_rows=np.nonzero((m2[:, :2] != m[:, :2]).any(1))[0])

    # the HDP spline tables themselves: parsed from the .nhdp by the PRODUCT's loader (sa_model_load; the reference
    # keeps them in structs private to hdp.c) and accepted only if the host density agrees with the reference's
    # get_nanopore_kmer_density on a dense probe set (every 97th kmer x 23 abscissae, incl. both extrapolation sides)
    import signalalign_b200 as sa
    ours = sa.Model.load(M_R73, NHDP)
    grid, k2r, pp, sl, own = ours.hdp_tables()
    worst = 0.0
    for ki in range(0, ours.K, 97):
        ks = ours.kmer_string(ki)
        for xq in np.linspace(-3.0, 103.0, 23):
            a, b = ours.hdp_density(ki, float(xq)), smh.hdp_density(ks, float(xq))
            worst = max(worst, abs(a - b))
    assert worst == 0.0, worst
    save("hdp_single_level_fixed", grid=grid, kmer_to_row=k2r, own_row=own, post_pred=pp, slopes=sl)

    # ------------------------------------------------------------------ R9.4 1-D read against its own sequence
    rd = ref.RefNpRead(R94_READ, "ACGT")
    seq = rd.template_read
    ev = rd.template_events()
    emap = rd.template_strand_event_map()
    sm = ref.RefStateMachine(M_R94, kind=0)
    rd.estimate_params(sm)   # signalUtils_estimateNanoporeParams: sets sM->scale/shift/var, drift-adjusts the events
    ev = rd.template_events()
    sc = sm.scaling()
    lX = len(seq) - 5
    # identity guide alignment -> anchors (nt i, event emap[i]) trimmed like a cigar block with constraintDiagonalTrim 14
    xs_ = np.arange(14, lX - 14)
    ys_ = emap[xs_]
    keep = np.concatenate([[True], np.diff(ys_) > 0])
    a94 = np.stack([xs_[keep], ys_[keep]], 1).astype(np.int64)
    cli = dict(threshold=0.01, min_diags=1000, trace_back=50, expansion=50, split_bigger=3000 * 3000)
    p, km, sec = sm.align(seq, ev, a94, **cli)
    print("r9p4 self-reference: %d events, %d kmers, %d anchors -> %d pairs in %.1f s" % (len(ev), lX, len(a94), len(p), sec))
    tr, lik, _, _, _ = sm.expectations(seq, ev, a94, **cli)
    save("r9p4_selfref", kind=np.int64(0), target=np.array(seq), events=ev, anchors=a94, scaling=sc,
         params=np.array(list(cli.values())), pairs=pairs4(p, km, sm.alphabet), exp_transitions=tr,
         exp_likelihood=np.float64(lik))

    # ------------------------------------------------------------------ unit-level known answers
    x = rng.uniform(-40, 5, 4000)
    y = x + rng.uniform(-9, 9, 4000)
    seams = np.array([0.0, 1.0, 2.5, 4.5, 7.5])
    eps = np.array([0.0, 1e-15, -1e-15, 1e-9, -1e-9])
    sx = rng.uniform(-30, 0, len(seams) * len(eps))
    sy = sx + (seams[:, None] + eps[None, :]).reshape(-1)
    x = np.concatenate([x, sx, sy, [-np.inf, -np.inf, 3.0]])
    y = np.concatenate([y, sy, sx, [-np.inf, 2.0, -np.inf]])
    la = np.array([ref.log_add(float(a), float(b)) for a, b in zip(x, y)])
    # emissions of the three Gaussian kinds on random (kmer, mean, noise) probes, match and gap-Y tables
    alpha = raw73.alphabet
    kk = ["".join(alpha[c] for c in rng.integers(0, 6, 6)) for _ in range(300)]
    em_mean = rng.uniform(40, 110, 300)
    em_noise = rng.uniform(0.3, 3.0, 300)
    em = {}
    for kind, tag in ((0, "meanonly"), (1, "scaled"), (2, "descale")):
        s = ref.RefStateMachine(M_R73, kind=kind)
        s.set_scaling(1.03, -1.7, 1.12)
        em["emit_" + tag] = np.array([[s.emission(k_, float(mu), float(nz), True), s.emission(k_, float(mu), float(nz), False)]
                                      for k_, mu, nz in zip(kk, em_mean, em_noise)])
    kid = np.array([ref.kmer_id(k_, alpha) for k_ in kk])
    # bands and split points on random anchor chains
    bands = {}
    for i in range(12):
        lX_, lY_ = int(rng.integers(1, 400)), int(rng.integers(1, 600))
        n = int(rng.integers(0, min(lX_, lY_) // 3 + 1))
        ax = np.sort(rng.choice(lX_, n, replace=False))
        ay = np.sort(rng.choice(lY_, n, replace=False))
        an = np.stack([ax, ay], 1).astype(np.int64)
        e = int(rng.choice([0, 2, 6, 20, 50]))
        bands["band%d_in" % i] = np.array([lX_, lY_, e])
        bands["band%d_anchors" % i] = an
        bands["band%d_out" % i] = ref.band(an, lX_, lY_, e)
        big = int(rng.choice([50, 400, 3000, 10 ** 7]))
        for rl in (0, 1):
            for rr in (0, 1):
                bands["split%d_%d%d" % (i, rl, rr)] = ref.split_points(an, lX_, lY_, big, rl, rr)
        bands["split%d_bigger" % i] = np.int64(big)
    save("unit", logadd_x=x, logadd_y=y, logadd_out=la, emit_kmers=np.array(kk), emit_mean=em_mean, emit_noise=em_noise,
         emit_scaling=np.array([1.03, -1.7, 1.12]), kmer_ids=kid, **em, **bands)


if __name__ == "__main__":
    if not ref.available():
        sys.exit("oracle/_ref/libsaref.so missing: run `make -C oracle ref` first")
    main()
