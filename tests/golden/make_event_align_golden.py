#!/usr/bin/env python
"""Golden vectors for SURVEY.md 8f N4 (adaptive banded event -> kmer alignment), generated HERE from the reference's own
adaptive_banded_simple_event_align2 / estimate_scalings_using_mom (impl/eventAligner.c, compiled unmodified into
oracle/_ref/libsaref_ea.so behind oracle/ea_harness.c).  Writes tests/golden/event_align.npz.
Needs /root/reference and `make -C oracle cli`."""
import ctypes as C, os, sys, tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(HERE))
from helpers import golden, write_model_file
from signalalign_b200 import synth

lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libsaref_ea.so"))
f32p, i64p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_double)
lib.ref_event_align.restype = C.c_int64
lib.ref_event_align.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int64, f32p, f32p, C.c_int, C.c_double, C.c_double, C.c_double,
                                C.c_int64, i64p, i64p, f64p]
m = golden("model_r94_acgt")
alphabet, k, table = str(m["alphabet"]), int(m["k"]), np.asarray(m["match"], np.float64).reshape(-1, 5)
rng = np.random.default_rng(20261106)
ref_codes = rng.integers(0, len(alphabet), 60000).astype(np.int64)
out = {"alphabet": np.array(alphabet), "k": np.array(k)}
cases = [dict(mean_len=1500, use_mom=1, rna=0), dict(mean_len=2500, use_mom=1, rna=0), dict(mean_len=1200, use_mom=0, rna=0),
         dict(mean_len=1000, use_mom=1, rna=1), dict(mean_len=900, use_mom=1, rna=0, garbage=True)]
with tempfile.TemporaryDirectory() as d:
    model = os.path.join(d, "r94.model")
    write_model_file(model, alphabet, k, m["trans"], m["match"])
    for ci, cs in enumerate(cases):
        r = synth.make_read(rng, ref_codes, table, alphabet, k, mean_len=cs["mean_len"], sigma_len=0.05, min_len=500, max_len=4000,
                            n_probs=(0.08, 0.62, 0.22, 0.08))
        mean = r["events"][:, 0].astype(np.float32)
        if cs["rna"]:
            # build_kmer_list (eventAligner.c:774-790) aligns to the REVERSED kmers (3'->5'): events for those
            seq = r["target"]
            kid = [sum(alphabet.index(c) * len(alphabet) ** (k - 1 - j) for j, c in enumerate(seq[i:i + k][::-1])) for i in range(len(seq) - k + 1)]
            n_ev = rng.choice(4, size=len(kid), p=(0.08, 0.62, 0.22, 0.08))
            own = np.repeat(np.arange(len(kid)), n_ev)
            mean = (table[np.array(kid)[own], 0] + 1.1 * table[np.array(kid)[own], 1] * rng.standard_normal(len(own))).astype(np.float32)
        if cs.get("garbage"):
            mean = rng.uniform(40, 140, len(mean)).astype(np.float32)       # no signal: the reference's QC rejects the alignment
        stdv = np.ones(len(mean), np.float32)
        seq = r["target"]
        cap = len(mean) + len(seq) + 8
        ki, ei, sc = np.zeros(cap, np.int64), np.zeros(cap, np.int64), np.zeros(3)
        n = lib.ref_event_align(model.encode(), seq.encode(), cs["rna"], len(mean), mean.ctypes.data_as(f32p), stdv.ctypes.data_as(f32p),
                                cs["use_mom"], r["scale"], r["shift"], r["var"], cap, ki.ctypes.data_as(i64p), ei.ctypes.data_as(i64p),
                                sc.ctypes.data_as(f64p))
        assert n >= 0
        print("case %d: %d nt, %d events, rna %d, mom %d -> %d aligned pairs, scalings %s" % (ci, len(seq), len(mean), cs["rna"], cs["use_mom"], n, sc))
        out["c%d_sequence" % ci] = np.array(seq)
        out["c%d_mean" % ci] = mean
        out["c%d_flags" % ci] = np.array([cs["rna"], cs["use_mom"]], np.int64)
        out["c%d_scalings" % ci] = sc
        out["c%d_pairs" % ci] = np.stack([ki[:n], ei[:n]], 1)
out["n_cases"] = np.array(len(cases))
out["match"] = np.asarray(m["match"], np.float64)
np.savez_compressed(os.path.join(HERE, "event_align.npz"), **out)
