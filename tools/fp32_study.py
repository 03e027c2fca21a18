#!/usr/bin/env python
"""FP32 question settled with data (CPU only): the oracle restatement built with
-DSAO_FP32_STUDY (float DP state relative to an integer per-diagonal offset, every logAdd / dot product / posterior in float:
what an FP32 band kernel with per-diagonal renormalisation would compute) against the exact FP64 restatement, on the golden
reads and three config-3 reads.  usage: make -C oracle port fp32study && python tools/fp32_study.py > profiles/r2_fp32_study.txt"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CHILD = r"""
import os, sys, pickle
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from oracle import port
if sys.argv[2] == "fp32":
    port.LIB_PATH = os.path.join(sys.argv[1], "oracle", "libsa_oracle_fp32.so")
from helpers import golden, golden_models, params_from
from signalalign_b200 import synth
out = {}
g = golden("zymo_meanonly"); tp = g["tparams"]
pm, _ = golden_models("model_r73_acegot", port.EMIT_GAUSS_MEANONLY_DESCALE)
pm.scale, pm.shift, pm.var = float(tp[0]), float(tp[1]), float(tp[2])
for tag in ("cli", "lib"):
    par = params_from(g["params_" + tag])
    out["zymo_" + tag] = port.align(pm, str(g["target"]), g["events"], g["anchors"], port.params(par["threshold"], par["min_diags"], par["trace_back"], par["expansion"], par["split_bigger"]))[0]
g = golden("r9p4_selfref")
pm, _ = golden_models("model_r94_acgt", port.EMIT_GAUSS_MEANONLY_DESCALE)
pm.scale, pm.shift, pm.var = [float(v) for v in g["scaling"]]
par = params_from(g["params"])
out["r9p4_selfref"] = port.align(pm, str(g["target"]), g["events"], g["anchors"], port.params(par["threshold"], par["min_diags"], par["trace_back"], par["expansion"], par["split_bigger"]))[0]
for pattern in ("dense", "contract"):
    reads, table = synth.make_batch(3, seed=20261104, **synth.ANCHOR_PATTERNS[pattern])
    pm = port.PortModel("ACGT", 6, synth.nanopore_transitions(), table)
    for i, r in enumerate(reads):
        pm.scale, pm.shift, pm.var = r["scale"], r["shift"], r["var"]
        out["config3_%s_%d" % (pattern, i)] = port.align(pm, r["target"], r["events"], r["anchors"], port.params(0.01, 1000, 50, 100))[0]
pickle.dump(out, open(sys.argv[3], "wb"))
"""


def main():
    import pickle
    import tempfile
    res = {}
    for mode in ("fp64", "fp32"):
        f = tempfile.mktemp(suffix=".pkl")
        subprocess.check_call([sys.executable, "-c", CHILD, ROOT, mode, f])
        res[mode] = pickle.load(open(f, "rb"))
    print("FP32 state + per-diagonal integer offset (oracle -DSAO_FP32_STUDY) vs the exact FP64 restatement; threshold 0.01")
    print("%-22s %8s %8s %10s %12s %14s %14s" % ("read", "pairs64", "pairs32", "only one", "max |dp|", "#|dp| > 1e-4", "#|dp| > 1e-5"))
    for k in res["fp64"]:
        a, b = res["fp64"][k], res["fp32"][k]
        da = {(int(x), int(y), int(km)): int(p) for p, x, y, km in a}
        db = {(int(x), int(y), int(km)): int(p) for p, x, y, km in b}
        common = set(da) & set(db)
        d = np.array([abs(da[q] - db[q]) for q in common]) / 1e7
        one = [max(da.get(q, 0), db.get(q, 0)) / 1e7 for q in set(da) ^ set(db)]
        print("%-22s %8d %8d %10s %12.2e %14d %14d" % (k, len(a), len(b), "%d (p<=%.4f)" % (len(one), max(one) if one else 0),
                                                        d.max(), int((d > 1e-4).sum()), int((d > 1e-5).sum())))


if __name__ == "__main__":
    main()
