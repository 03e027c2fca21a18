#!/usr/bin/env python
"""Golden vectors for the per-read scaling estimation from the COMPILED REFERENCE
(oracle/_ref/libsaref.so: signalUtils_estimateNanoporeParams, impl/signalMachineUtils.c:186-225) on the two .npRead
fixtures of the reference's tests.  Runs where /root/reference exists:  python tests/golden/make_estimate_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref  # noqa: E402

REF = os.environ.get("SA_REFERENCE", "/root/reference")
CASES = {
    "r9p4": ("tests/test_npReads/r9p4_oneD.npRead", "models/testModelR9.4_450bps.nucleotide.6mer.template.model", "ACGT", "model_r94_acgt"),
    "zymo": ("tests/test_npReads/ZymoC_ch_1_file1.npRead", "models/testModelR73_acegot_template.model", "ACEGOT", "model_r73_acegot"),
}
NAMES = ["scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift"]


def main():
    out = {}
    for tag, (npfile, model, alphabet, table) in CASES.items():
        sm = ref.RefStateMachine(os.path.join(REF, model), kind=0)
        r = ref.RefNpRead(os.path.join(REF, npfile), alphabet)
        out[tag + "_read"] = np.array(r.template_read)
        out[tag + "_map"] = r.template_strand_event_map()
        out[tag + "_events"] = r.template_events()
        out[tag + "_table"] = np.array(table)
        r.estimate_params(sm)
        out[tag + "_params"] = np.array([r.template_params[k] for k in NAMES])
        out[tag + "_means_after"] = r.template_events()[:, 0].copy()
        print(tag, dict(zip(NAMES, out[tag + "_params"])))
    np.savez_compressed(os.path.join(HERE, "estimate_params.npz"), **out)


if __name__ == "__main__":
    main()
