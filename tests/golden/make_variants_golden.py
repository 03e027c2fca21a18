#!/usr/bin/env python
"""Golden vectors for SURVEY.md 8f N3 (variant calls -> per-position marginals), generated HERE from the reference:
  * rows: the reference's own writePosteriorProbsVC (impl/signalMachine.c:161-232, compiled unmodified into
    oracle/_ref/libsaref_vc.so, flat entry oracle/vc_harness.c) on synthetic aligned pairs, all four
    (strand, mapping) combinations;
  * marginals: the reference's own MarginalizeVariants class (src/signalalign/variantCaller.py:18-89), its source
    extracted with ast and executed with stand-ins for the two helpers it imports (py3helpers.merge_lists,
    NanoporeRead.bytes_to_string) — the package itself cannot be imported here (py3helpers, h5py absent) — on the TSV
    the writer produced, read the way SignalAlignment.read_in_signal_align_tsv reads it (signalAlignment.py:652-668).
Writes tests/golden/variants.npz.  Needs /root/reference and `make -C oracle cli`."""
import ast, ctypes as C, itertools, os, sys, tempfile
import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libsaref_vc.so"))
i64p = C.POINTER(C.c_int64)
lib.ref_write_vc.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, i64p, i64p, i64p,
                             C.c_char_p, C.c_int, C.c_double, C.c_int, C.c_char_p]

src = open(os.path.join(REF, "src", "signalalign", "variantCaller.py")).read()
cls = [n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == "MarginalizeVariants"][0]
ns = {"pd": pd, "np": np, "merge_lists": lambda ls: list(itertools.chain.from_iterable(ls)),
      "NanoporeRead": type("NanoporeRead", (), {"bytes_to_string": staticmethod(lambda b: b.decode() if isinstance(b, bytes) else str(b))})}
exec(compile(ast.Module(body=[cls], type_ignores=[]), "variantCaller.py", "exec"), ns)
MarginalizeVariants = ns["MarginalizeVariants"]

VC_DTYPE = [('event_index', int), ('reference_position', int), ('base', 'S6'), ('posterior_probability', float), ('strand', 'S1'),
            ('forward_mapped', "S8"), ('read_file', 'S57'), ('posterior_score', float), ('contig', 'S100')]

rng = np.random.default_rng(20261105)
k, L = 6, 1200
out = {}
cases = []
for ci, (variants, comp, fwd) in enumerate([("CE", 0, 1), ("CE", 0, 0), ("ACGT", 1, 1), ("CEO", 1, 0)]):
    t = rng.choice(list("ACGT"), L)
    xs = rng.choice(np.arange(20, L - 20), 60, replace=False)
    t[xs] = "X"
    target = "".join(t)
    pairs, kmers = [], []
    for x in range(0, L - k + 1):
        for _ in range(int(rng.integers(0, 4))):
            kmer = "".join(c if c != "X" else rng.choice(list(variants)) for c in target[x:x + k])
            p = int(rng.choice([rng.integers(100000, 10000001), rng.integers(100000, 200000), 10 * rng.integers(10000, 1000000) + 5]))
            pairs.append((p, x, int(rng.integers(0, 4000)), 0))
            kmers.append(kmer)
    pairs = np.array(pairs, dtype=np.int64)
    ev_off, ref_off = 17 + ci, 1000 * ci
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "vc.tsv")
        c = lambda a: np.ascontiguousarray(a).ctypes.data_as(i64p)
        pcol, xcol, ycol = (np.ascontiguousarray(pairs[:, j]) for j in range(3))
        lib.ref_write_vc(path.encode(), b"read1", k, target.encode(), fwd, ev_off, ref_off, len(pairs), c(pcol), c(xcol), c(ycol),
                         "".join(kmers).encode(), comp, 12.5, 0, b"contig1")
        text = open(path).read()
        table = np.loadtxt(path, dtype=VC_DTYPE, ndmin=1)
    names = [n for n in table.dtype.names if n != "read_file"]
    mv = MarginalizeVariants(table[names], variants=list(variants), read_name="read1")
    pos = mv.get_data()
    pos = pos.sort_values("position")
    out["c%d_target" % ci] = np.array(target)
    out["c%d_pairs" % ci] = pairs
    out["c%d_kmers" % ci] = np.array(kmers)
    out["c%d_meta" % ci] = np.array([k, comp, fwd, ev_off, ref_off], dtype=np.int64)
    out["c%d_variants" % ci] = np.array(variants)
    out["c%d_tsv" % ci] = np.array(text)
    out["c%d_positions" % ci] = pos["position"].to_numpy(dtype=np.int64)
    out["c%d_probs" % ci] = pos[sorted(variants)].to_numpy(dtype=np.float64)
    out["c%d_per_read" % ci] = mv.per_read_calls[sorted(variants)].to_numpy(dtype=np.float64)[0]
    print("case %d: %d pairs, %d rows, %d positions" % (ci, len(pairs), len(table), len(pos)))
out["n_cases"] = np.array(4)
np.savez_compressed(os.path.join(HERE, "variants.npz"), **out)
