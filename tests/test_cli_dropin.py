"""The drop-in at the CLI: the reference's REAL `signalMachine` (impl/signalMachine.c, compiled
unmodified) linked twice by `make -C oracle cli` — with the reference's own objects (signalMachine_cpu) and with the shim
+ libsagpu.so in place of pairwiseAligner.c's two entry points (signalMachine_gpu) — must write the same `full`,
`assignments` and `variantCaller` TSVs (impl/signalMachine.c:89-270) and the same expectations file
(impl/continuousHmm.c:353-407): every column byte-equal, posteriors within 1e-4, rows present on one side only must sit
within 1e-4 of the threshold.  The golden files were written by signalMachine_cpu (tests/golden/make_cli_golden.py).

The binaries are built where /root/reference exists and travel to the GPU box with the snapshot (oracle/_ref/).
"""
import os
import subprocess
import tempfile

import pytest

import cli_cases as cc

CPU_BIN = os.path.join(cc.REF_BIN, "signalMachine_cpu")
GPU_BIN = os.path.join(cc.REF_BIN, "signalMachine_gpu")
P_COL = {"full": [12], "assignments": [3], "variantCaller_default_table": [3, 7], "variantCaller_custom_table": [3, 7]}
THRESHOLD, TOL = 0.01, 1e-4


def _run(binary, case, work, inp):
    out = os.path.join(work, case + ".out")
    if os.path.exists(out):
        os.remove(out)
    r = subprocess.run(cc.command(binary, inp, case, out), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SUCCESS" in r.stderr, (case, r.stderr[-800:])   # what signalAlignment.py:480 looks for
    return open(out).read(), r.stdout


def _compare_tsv(case, got, want):
    cols = P_COL[case]

    def key(fields):
        return tuple(f for i, f in enumerate(fields) if i not in cols)
    g = [l.split("\t") for l in got.splitlines()]
    w = [l.split("\t") for l in want.splitlines()]
    if [key(f) for f in g] == [key(f) for f in w]:      # same rows in the same order: posteriors pairwise
        for a, b in zip(g, w):
            for c in cols:
                assert abs(float(a[c]) - float(b[c])) <= TOL + 1e-6, (case, a, b)
        return
    # otherwise the difference may only be rows within tolerance of the threshold (a variantCaller row is not unique by
    # its other columns — several kmer windows report the same position for the same event — hence multisets)
    import collections
    cg, cw = collections.Counter(key(f) for f in g), collections.Counter(key(f) for f in w)
    extra = (cg - cw) + (cw - cg)
    near = collections.Counter(key(f) for f in g + w if float(f[cols[0]]) <= THRESHOLD + TOL + 1e-6)
    for k, n in extra.items():
        assert near[k] >= n, (case, "row on one side only and not near the threshold", k)
    assert sum(extra.values()) <= 4, (case, extra)


def _compare_expectations(got, want):
    gl, wl = got.splitlines(), want.splitlines()
    assert len(gl) == len(wl)
    for i, (a, b) in enumerate(zip(gl, wl)):
        ta, tb = a.split(), b.split()
        assert len(ta) == len(tb), i
        for x, y in zip(ta, tb):
            try:
                fx, fy = float(x), float(y)
            except ValueError:
                assert x == y, (i, x, y)
                continue
            assert abs(fx - fy) <= 1e-5 * max(abs(fy), 1e-3), (i, x, y)


@pytest.mark.skipif(not os.path.exists(CPU_BIN), reason="oracle/_ref/signalMachine_cpu not built (needs /root/reference)")
def test_reference_cli_reproduces_the_golden_files():
    with tempfile.TemporaryDirectory(prefix="sa_cli_") as work:
        inp = cc.prepare_inputs(work)
        for case, (_, kind) in cc.CASES.items():
            got, _ = _run(CPU_BIN, case, work, inp)
            assert got == cc.read_golden(case), case


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(GPU_BIN), reason="oracle/_ref/signalMachine_gpu not built (needs /root/reference)")
def test_signalmachine_linked_against_libsagpu_writes_the_same_files():
    with tempfile.TemporaryDirectory(prefix="sa_cli_") as work:
        inp = cc.prepare_inputs(work)
        for case, (_, kind) in cc.CASES.items():
            got, stdout = _run(GPU_BIN, case, work, inp)
            want = cc.read_golden(case)
            if kind == "expectations":
                _compare_expectations(got, want)
            else:
                _compare_tsv(case, got, want)
                assert stdout.startswith("r9read 6514\t")        # "<label> <anchors>\t<pairs>(<score>)", signalMachine.c:917
        # a custom ambiguity table must change the calls (the shim once dropped it and used the built-in table)
        assert cc.read_golden("variantCaller_custom_table") != cc.read_golden("variantCaller_default_table")


BATCH_BIN = os.path.join(cc.REF_BIN, "signalMachineBatch")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(BATCH_BIN), reason="oracle/_ref/signalMachineBatch not built (needs /root/reference)")
def test_signalmachinebatch_manifest_writes_the_same_files():
    """one process, model parsed once, all reads of the manifest in one sa_gpu_align_batch call, every read's
    file written by the reference's own writers — same bytes as the per-read CLI (golden from signalMachine_cpu)."""
    with tempfile.TemporaryDirectory(prefix="sa_cli_") as work:
        inp = cc.prepare_inputs(work)

        def run(fmt, entries, extra=()):
            man = os.path.join(work, "manifest_%s.tsv" % fmt)
            outs = []
            with open(man, "w") as f:
                f.write("# npRead\tcigar\tfasta\tsequence\tlabel\toutput\n")
                for i, fasta in enumerate(entries):
                    out = os.path.join(work, "batch_%s_%d.tsv" % (fmt, i))
                    if os.path.exists(out):
                        os.remove(out)
                    outs.append(out)
                    f.write("\t".join([inp["npread"], inp["cigar"], fasta, "r9ref", "r9read", out]) + "\n")
            cmd = [BATCH_BIN, "--manifest", man, "-T", inp["model"], "-s", str(fmt), "-x", "50", "-D", "0.01", "-m", "14", "--batch", "2"]
            r = subprocess.run(cmd + list(extra), capture_output=True, text=True, timeout=900)
            assert r.returncode == 0, r.stderr[-800:]
            assert r.stderr.count("SUCCESS") == len(entries) and len(r.stdout.strip().splitlines()) == len(entries)
            return [open(o).read() for o in outs], r.stdout

        outs, stdout = run(0, [inp["fasta"], inp["fasta"], inp["fasta"]])      # two batches (2 + 1 reads)
        for o in outs:
            _compare_tsv("full", o, cc.read_golden("full"))
        assert all(l.startswith("r9read 6514\t") for l in stdout.strip().splitlines())
        outs, _ = run(2, [inp["fasta"]])
        _compare_tsv("assignments", outs[0], cc.read_golden("assignments"))
        # the scaling re-estimation runs on the device by default (bit-identical): the reference's host function gives the
        # same bytes
        host, _ = run(0, [inp["fasta"]], extra=["--host-estimate"])
        assert host[0] == run(0, [inp["fasta"]])[0][0]
        outs, _ = run(1, [inp["fasta_variants"], inp["fasta"]])
        _compare_tsv("variantCaller_default_table", outs[0], cc.read_golden("variantCaller_default_table"))
        assert outs[1] == ""                                                      # no ambiguity code: nothing to report
        outs, _ = run(1, [inp["fasta_variants"]], extra=["-a", inp["ambig"]])
        _compare_tsv("variantCaller_custom_table", outs[0], cc.read_golden("variantCaller_custom_table"))
