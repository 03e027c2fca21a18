"""-m gpu: BASELINE config 5 AT SIZE — one 5-mer (K = 1024) read of more than 50 000 events, Gaussian and HDP emission
(synthetic spline tables in the shape of the absent RNA .nhdp: one row per kmer, G = 1200 grid points), against the oracle,
in every way the scheduler can run a long read: split mode (forward-only sweep + independent traceback blocks: the default
for a read that dominates its batch), the latency class in ring mode, the throughput mapping, and the wide kernel — all
byte-identical to each other."""
import os

import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port
from signalalign_b200 import synth

from helpers import compare_pairs, oracle_align, port_model_from, port_params, sa_read

pytestmark = pytest.mark.gpu
N_PROBS = (0.10, 0.55, 0.25, 0.10)
MODES = {"split": {}, "latency_ring": {"SA_NO_SPLIT": "1"}, "throughput": {"SA_NO_LATENCY_CLASS": "1"},
         "wide": {"SA_WIDE_ONLY": "1"}}


def _long_read(pattern):
    table = synth.gaussian_table("ACGT", 5, seed=9)
    reads, _ = synth.make_batch(1, seed=77, alphabet="ACGT", k=5, table=table, mean_len=40000, sigma_len=0.01, min_len=39000,
                                max_len=41000, n_probs=N_PROBS, **synth.ANCHOR_PATTERNS[pattern])
    assert reads[0]["events"].shape[0] > 50000
    return reads[0], table


def _hdp_tables(table, G=1200):
    grid = np.linspace(50.0, 140.0, G)
    mu = table[:, 0]
    pp = 0.7 * np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / 2.0) ** 2) / (2.0 * np.sqrt(2 * np.pi)) + \
        0.3 * np.exp(-0.5 * ((grid[None, :] - mu[:, None] - 3.0) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
    return grid, np.arange(table.shape[0], dtype=np.int32), pp, np.gradient(pp, grid, axis=1)


def _run(model, read, p, env):
    for k, v in env.items():
        os.environ[k] = v
    try:
        b = sa.Batch(model, [sa_read(read)], p)
        kernels = b.kernel_items
        b.run()
        res = b.fetch()
        b.close()
    finally:
        for k in env:
            os.environ.pop(k, None)
    return res.read_pairs(0).copy(), kernels


@pytest.mark.parametrize("hdp", [False, True])
def test_long_5mer_read_every_scheduling_mode(hdp):
    read, table = _long_read("dense")
    trans = synth.nanopore_transitions()
    p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
    if hdp:
        grid, k2r, pp, slopes = _hdp_tables(table)
        model = sa.Model.from_tables("ACGT", 5, trans, table, kind=sa.EMIT_HDP)
        model.set_hdp(grid, k2r, pp, slopes)
        pm = port_model_from("ACGT", 5, trans, table, kind=port.EMIT_HDP, grid=grid, kmer_to_row=k2r, post_pred=pp, slopes=slopes)
    else:
        model = sa.Model.from_tables("ACGT", 5, trans, table)
        pm = port_model_from("ACGT", 5, trans, table)
    ora, _ = oracle_align(pm, read, port_params(p))
    first = None
    for name, env in MODES.items():
        got, kernels = _run(model, read, p, env)
        assert kernels[0] == 0 and (kernels[3] == 1) == (name == "wide"), (name, kernels)
        assert compare_pairs(got, ora, p.threshold) <= 20, name
        if first is None:
            first = got
        assert np.array_equal(got, first), name                  # the modes agree to the last bit


def test_long_5mer_read_contract_anchors_wide_kernel_and_expectations():
    """the same length with sparse (trim-14) anchors: bands far wider than 256 -> the wide kernel by default"""
    read, table = _long_read("contract")
    trans = synth.nanopore_transitions()
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=1000)
    model = sa.Model.from_tables("ACGT", 5, trans, table)
    pm = port_model_from("ACGT", 5, trans, table)
    got, kernels = _run(model, read, p, {})
    assert kernels == (0, 0, 0, 1), kernels
    ora, _ = oracle_align(pm, read, port_params(p))
    assert compare_pairs(got, ora, p.threshold) <= 20
    pm.scale, pm.shift, pm.var = read["scale"], read["shift"], read["var"]
    t, lik = port.expectations(pm, read["target"], read["events"], read["anchors"], port_params(p), pseudo_t=0.0)[:2]
    e = sa.expectations_batch(model, [sa_read(read)], p)
    assert np.allclose(e.transitions[0], t, rtol=1e-6, atol=1e-9) and abs(e.likelihood[0] - lik) <= 1e-5 * abs(lik)
