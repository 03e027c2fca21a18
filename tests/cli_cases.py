"""The signalMachine command lines of the CLI drop-in test (tests/test_cli_dropin.py) and of its golden generator
(tests/golden/make_cli_golden.py): the reference's REAL CLI (impl/signalMachine.c, compiled unmodified by
`make -C oracle cli`) on the R9.4 1-D fixture, once per output format plus the expectations pass.

Inputs are rebuilt from committed fixtures so that the same bytes exist here and on the GPU box:
  * tests/golden/cli/r9p4_oneD.npRead.gz — the reference's own test read (tests/test_npReads/r9p4_oneD.npRead, a data
    fixture), 10 922 events; the E. coli reference it was aligned to is missing from the snapshot, so the read's own
    base-called sequence (line 3 of the .npRead) is the reference and the guide alignment is the identity cigar;
  * the .model file is written from tests/golden/model_r94_acgt.npz by helpers.write_model_file.
"""
import gzip
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CLI_DIR = os.path.join(HERE, "golden", "cli")
REF_BIN = os.path.join(os.path.dirname(HERE), "oracle", "_ref")

# name -> (extra arguments, output kind)
CASES = {
    "full": (["-s", "0"], "posteriors"),
    "assignments": (["-s", "2"], "posteriors"),
    "variantCaller_default_table": (["-s", "1", "--variants"], "posteriors"),
    "variantCaller_custom_table": (["-s", "1", "--variants", "--ambig"], "posteriors"),
    "expectations": (["--expectations"], "expectations"),
}


def prepare_inputs(workdir):
    """writes the .npRead, FASTA files, cigar, ambiguity file and .model into workdir; returns their paths"""
    from helpers import golden, write_model_file
    os.makedirs(workdir, exist_ok=True)
    np_path = os.path.join(workdir, "r9p4_oneD.npRead")
    with gzip.open(os.path.join(CLI_DIR, "r9p4_oneD.npRead.gz"), "rb") as f, open(np_path, "wb") as o:
        o.write(f.read())
    seq = open(np_path).read().split("\n")[2].strip()
    fa = os.path.join(workdir, "ref.fa")
    with open(fa, "w") as f:
        f.write(">r9ref\n" + seq + "\n")
    # variant reference: every 97th position becomes the ambiguity code X
    v = list(seq)
    for i in range(50, len(v) - 10, 97):
        v[i] = "X"
    fav = os.path.join(workdir, "ref_variants.fa")
    with open(fav, "w") as f:
        f.write(">r9ref\n" + "".join(v) + "\n")
    cigar = os.path.join(workdir, "guide.cigar")
    with open(cigar, "w") as f:
        f.write("cigar: r9read 0 %d + r9ref 0 %d + 1000 M %d\n" % (len(seq), len(seq), len(seq)))
    ambig = os.path.join(workdir, "ambig.tsv")
    with open(ambig, "w") as f:        # `code<TAB>bases` per line (create_ambig_bases2, impl/pairwiseAligner.c:68-92)
        f.write("X\tCT\nR\tAG\n")
    m = golden("model_r94_acgt")
    model = os.path.join(workdir, "r94.model")
    write_model_file(model, str(m["alphabet"]), int(m["k"]), m["trans"], m["match"])
    return dict(npread=np_path, fasta=fa, fasta_variants=fav, cigar=cigar, ambig=ambig, model=model)


def command(binary, inp, case, out_path):
    extra, kind = CASES[case]
    cmd = [binary, "-q", inp["npread"], "-p", inp["cigar"], "-n", "r9ref", "-T", inp["model"], "-L", "r9read",
           "-x", "50", "-D", "0.01", "-m", "14"]
    fasta = inp["fasta"]
    args = []
    for a in extra:
        if a == "--variants":
            fasta = inp["fasta_variants"]
        elif a == "--ambig":
            args += ["-a", inp["ambig"]]
        elif a == "--expectations":
            pass
        else:
            args.append(a)
    cmd += ["-f", fasta] + args
    cmd += ["-t", out_path] if kind == "expectations" else ["-u", out_path]
    return cmd


def golden_path(case):
    return os.path.join(CLI_DIR, case + ".out.gz")


def read_golden(case):
    with gzip.open(golden_path(case), "rt") as f:
        return f.read()
