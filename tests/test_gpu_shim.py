"""-m gpu: the reference's OWN entry points on the GPU.

oracle/_ref/libsaref_gpu.so is the unmodified reference (compiled by oracle/Makefile) with getAlignedPairsUsingAnchors /
getExpectationsUsingAnchors bound to libsagpu.so by signalalign_b200/shim/signalalign_shim.c — the link-time
substitution INTEGRATION.md describes.  The reference's loaders build the StateMachine from a .model file, the harness
calls the entry points exactly as signalMachine.c does, and the results are compared with the golden outputs of the
unmodified reference (tests/golden/).  Runs in a process of its own: two builds of the same C library must not share a
symbol namespace.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from helpers import (ROOT_DIR, compare_expectations, compare_pairs, golden, params_from, write_golden_hdp_files,
                     write_model_file)

pytestmark = pytest.mark.gpu

GPU_LIB = os.path.join(ROOT_DIR, "oracle", "_ref", "libsaref_gpu.so")

CHILD = r"""
import ctypes, json, os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from oracle import ref
job = json.load(open(sys.argv[2]))
z = np.load(job["inputs"])
sm = ref.RefStateMachine(job["model"], kind=0, nhdp_file=job.get("nhdp"))
sm.set_scaling(*job["scaling"])
out = {"sm_type": np.array(sm.type)}
for tag, par in job["params"].items():
    pairs, kmers, sec = sm.align(str(z["target"]), z["events"], z["anchors"], **par)
    out["pairs_" + tag] = pairs
    out["kmers_" + tag] = np.array(kmers)
    epar = dict(par, **job.get("exp_params", {}).get(tag, {}))
    tr, lik, ak, ae, _ = sm.expectations(str(z["target"]), z["events"], z["anchors"], **epar)
    out["tr_" + tag] = tr
    out["lik_" + tag] = np.array(lik)
    out["ak_" + tag] = ak
    out["ae_" + tag] = ae
sagpu = ctypes.CDLL(os.path.join(sys.argv[1], "signalalign_b200", "libsagpu.so"))
a, b, c = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
sagpu.sa_last_call_stats(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
out["gpu_launches_last_align"] = np.array(c.value)
out["has_cpu_symbol"] = np.array(hasattr(ref.lib(), "cpu_getAlignedPairsUsingAnchors"))
np.savez(job["out"], **out)
"""


def _run_child(tmp_path, model_file, scaling, inputs, params, nhdp_file=None, exp_params=None):
    job = dict(model=model_file, scaling=[float(v) for v in scaling], inputs=os.path.join(str(tmp_path), "in.npz"),
               params=params, out=os.path.join(str(tmp_path), "out.npz"), nhdp=nhdp_file, exp_params=exp_params or {})
    np.savez(job["inputs"], **inputs)
    with open(os.path.join(str(tmp_path), "job.json"), "w") as f:
        json.dump(job, f)
    env = dict(os.environ, SA_REF_LIB=GPU_LIB)
    r = subprocess.run([sys.executable, "-c", CHILD, ROOT_DIR, os.path.join(str(tmp_path), "job.json")], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    with np.load(job["out"]) as z:
        return {k: z[k] for k in z.files}


def _kw(par):
    return dict(threshold=par["threshold"], min_diags=par["min_diags"], trace_back=par["trace_back"],
                expansion=par["expansion"], split_bigger=par["split_bigger"])


@pytest.mark.skipif(not os.path.exists(GPU_LIB), reason="oracle/_ref/libsaref_gpu.so not built (needs /root/reference at build time)")
def test_reference_entry_points_run_on_the_gpu(tmp_path):
    g = golden("zymo_meanonly")
    t = golden("model_r73_acegot")
    model_file = os.path.join(str(tmp_path), "acegot.model")
    write_model_file(model_file, str(t["alphabet"]), int(t["k"]), t["trans"], t["match"])
    tags = ("cli", "lib", "tight", "split")
    params = {tag: _kw(params_from(g["params_" + tag])) for tag in tags}
    out = _run_child(tmp_path, model_file, g["tparams"][:3], dict(target=g["target"], events=g["events"], anchors=g["anchors"]),
                     params)
    assert bool(out["has_cpu_symbol"]) and int(out["gpu_launches_last_align"]) > 0      # the shim build, on the device
    alphabet, k = str(t["alphabet"]), int(t["k"])
    for tag in tags:
        want = g["pairs_" + tag]                      # (p, x, y, kmer index) from the unmodified reference
        got = out["pairs_" + tag]
        idx = np.array([sum(alphabet.index(c) * len(alphabet) ** (k - 1 - i) for i, c in enumerate(s)) for s in out["kmers_" + tag]],
                       np.int64)
        got4 = np.concatenate([got, idx[:, None]], axis=1)
        # the reference returns its list unordered per block; compare as the other GPU tests do (same order expected)
        assert compare_pairs(got4.astype(np.int32), want, params[tag]["threshold"]) <= 20, tag
        compare_expectations(out["tr_" + tag], float(out["lik_" + tag]), g["exp_transitions_" + tag],
                             float(g["exp_likelihood_" + tag]))


@pytest.mark.skipif(not os.path.exists(GPU_LIB), reason="oracle/_ref/libsaref_gpu.so not built (needs /root/reference at build time)")
def test_reference_hdp_machine_runs_on_the_gpu(tmp_path):
    """StateMachine3_HDP through the struct: the reference's deserialize_nhdp builds the machine, the shim hands its
    NanoporeHDP to libsagpu through serialize_nhdp, and pairs, transition sums and the (kmer, event) assignment list
    match the unmodified reference (1217 pairs at 0.1: tests/stateMachineTests.c:902-918)."""
    g = golden("zymo_hdp")
    model_file, nhdp_file = write_golden_hdp_files(str(tmp_path))
    par = _kw(params_from(g["params"]))
    exp_thr = 0.01                                     # the golden expectations ran with the library defaults
    out = _run_child(tmp_path, model_file, g["tparams"][:3], dict(target=g["target"], events=g["events"], anchors=g["anchors"]),
                     dict(hdp=par), nhdp_file=nhdp_file, exp_params=dict(hdp=dict(threshold=exp_thr)))
    assert bool(out["has_cpu_symbol"]) and int(out["gpu_launches_last_align"]) > 0
    assert int(out["sm_type"]) == 7                    # threeStateHdp (inc/stateMachine.h:37-46)
    t = golden("model_r73_acegot")
    alphabet, k = str(t["alphabet"]), int(t["k"])
    idx = np.array([sum(alphabet.index(c) * len(alphabet) ** (k - 1 - i) for i, c in enumerate(s)) for s in out["kmers_hdp"]], np.int64)
    got4 = np.concatenate([out["pairs_hdp"], idx[:, None]], axis=1)
    assert len(got4) == 1217
    assert compare_pairs(got4.astype(np.int32), g["pairs"], par["threshold"]) <= 20
    compare_expectations(out["tr_hdp"], float(out["lik_hdp"]), g["exp_transitions"], float(g["exp_likelihood"]))
    assert np.array_equal(out["ak_hdp"], g["exp_assign_kmer"]) and np.array_equal(out["ae_hdp"], g["exp_assign_event"])


MUTATE_CHILD = r"""
import json, os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from oracle import ref
job = json.load(open(sys.argv[2]))
z = np.load(job["inputs"])
sm = ref.RefStateMachine(job["model"], kind=2)
sm.set_scaling(1.0, 0.0, 1.0)
first, _, _ = sm.align(str(z["target"]), z["events"], z["anchors"], **job["params"])
sm.scale_model(*job["tparams"])              # emissions_signal_scaleModel: rewrites the tables IN PLACE
second, kmers, _ = sm.align(str(z["target"]), z["events"], z["anchors"], **job["params"])
np.savez(job["out"], first=first, second=second, kmers=np.array(kmers))
"""


@pytest.mark.skipif(not os.path.exists(GPU_LIB), reason="oracle/_ref/libsaref_gpu.so not built (needs /root/reference at build time)")
def test_machine_mutated_in_place_between_calls_is_uploaded_again(tmp_path):
    """the shim caches device models by CONTENT: after emissions_signal_scaleModel rewrote the tables of the same
    StateMachine in place, the next call must see the new tables (golden: loadScaledStateMachine3's 1076 pairs)"""
    g = golden("zymo_scaled")
    t = golden("model_r73_acegot")
    model_file = os.path.join(str(tmp_path), "acegot.model")
    write_model_file(model_file, str(t["alphabet"]), int(t["k"]), t["trans"], t["match"])
    tp = [float(v) for v in g["tparams"]]
    job = dict(model=model_file, inputs=os.path.join(str(tmp_path), "in.npz"), out=os.path.join(str(tmp_path), "out.npz"),
               params=_kw(params_from(g["params"])), tparams=[tp[0], tp[1], tp[2], tp[4], tp[5]])
    np.savez(job["inputs"], target=g["target"], events=g["events"], anchors=g["anchors"])
    with open(os.path.join(str(tmp_path), "job.json"), "w") as f:
        json.dump(job, f)
    r = subprocess.run([sys.executable, "-c", MUTATE_CHILD, ROOT_DIR, os.path.join(str(tmp_path), "job.json")],
                       env=dict(os.environ, SA_REF_LIB=GPU_LIB), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    with np.load(job["out"]) as z:
        first, second, kmers = z["first"], z["second"], z["kmers"]
    alphabet, k = str(t["alphabet"]), int(t["k"])
    idx = np.array([sum(alphabet.index(c) * len(alphabet) ** (k - 1 - i) for i, c in enumerate(s)) for s in kmers], np.int64)
    got4 = np.concatenate([second, idx[:, None]], axis=1)
    assert len(got4) == 1076 and not np.array_equal(first, second)
    assert compare_pairs(got4.astype(np.int32), g["pairs"], job["params"]["threshold"]) <= 20
