#!/usr/bin/env python
"""Golden outputs of the reference's real CLI (oracle/_ref/signalMachine_cpu = impl/signalMachine.c + the unmodified
reference objects, `make -C oracle cli`) for tests/test_cli_dropin.py.  Runs where /root/reference exists.

usage: python tests/golden/make_cli_golden.py
"""
import gzip
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cli_cases as cc  # noqa: E402

REF = os.environ.get("SA_REFERENCE", "/root/reference")


def main():
    os.makedirs(cc.CLI_DIR, exist_ok=True)
    src = os.path.join(REF, "tests", "test_npReads", "r9p4_oneD.npRead")
    with open(src, "rb") as f, gzip.GzipFile(os.path.join(cc.CLI_DIR, "r9p4_oneD.npRead.gz"), "wb", mtime=0) as o:
        o.write(f.read())
    binary = os.path.join(cc.REF_BIN, "signalMachine_cpu")
    work = tempfile.mkdtemp(prefix="sa_cli_golden_")
    inp = cc.prepare_inputs(work)
    for case in cc.CASES:
        out = os.path.join(work, case + ".out")
        cmd = cc.command(binary, inp, case, out)
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0 and "SUCCESS" in r.stderr, (case, r.stderr[-500:])
        data = open(out, "rb").read()
        with gzip.GzipFile(cc.golden_path(case), "wb", mtime=0) as o:
            o.write(data)
        print(case, len(data), "bytes,", data.count(b"\n"), "lines;", r.stdout.strip())
    shutil.rmtree(work)


if __name__ == "__main__":
    main()
