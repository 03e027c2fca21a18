"""Per-read scale / shift / var / drift estimation (reference signalUtils_estimateNanoporeParams,
impl/signalMachineUtils.c:186-225): the oracle port against golden vectors of the compiled reference (CPU), and the device
kernel through the C-ABI (sa_gpu_estimate_params_batch) against the golden vectors and against the oracle on synthetic
strands.  Every sum keeps the reference's order, so the bar is BIT-exact parameters and event means."""
import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port
from signalalign_b200 import synth

from helpers import golden

NAMES = ["scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift"]


def _case(tag):
    g = golden("estimate_params")
    m = golden(str(g[tag + "_table"]))
    return g, m, str(g[tag + "_read"]), g[tag + "_map"], g[tag + "_events"]


@pytest.mark.parametrize("tag", ["r9p4", "zymo"])
def test_oracle_port_reproduces_the_reference(tag):
    g, m, read, emap, events = _case(tag)
    got, ev, n = port.estimate_params(m["match"], str(m["alphabet"]), int(m["k"]), read, emap, events)
    assert n > 0
    assert [got[k] for k in NAMES] == list(g[tag + "_params"])           # bit-exact
    assert np.array_equal(ev[:, 0], g[tag + "_means_after"])
    assert np.array_equal(ev[:, 1:], events[:, 1:])


def _synthetic_strands(n, seed=3):
    rng = np.random.default_rng(seed)
    table = synth.gaussian_table()
    out = []
    for i in range(n):
        L = int(rng.integers(40, 3000))
        codes = rng.integers(0, 4, L)
        read = "".join("ACGT"[c] for c in codes)
        steps = rng.choice([0, 1, 1, 1, 2, 3], L)            # stays, skips: the map is non-decreasing with repeats
        emap = np.cumsum(steps) - steps[0]
        n_ev = int(emap[-1]) + 1 + int(rng.integers(0, 5))
        kidx = synth.kmer_indices(codes, 6, 4)
        ev = np.zeros((n_ev, 4))
        ev[:, 0] = rng.uniform(60, 125, n_ev)
        own = np.searchsorted(emap[:len(kidx)], np.arange(n_ev), side="left").clip(0, len(kidx) - 1)
        ev[:, 0] = 1.03 * table[kidx[own], 0] + 2.0 + rng.standard_normal(n_ev) * 1.5
        ev[:, 1] = rng.uniform(0.5, 2.5, n_ev)
        ev[:, 2] = 0.002
        ev[:, 3] = np.cumsum(rng.uniform(0.001, 0.01, n_ev))
        ev[:, 0] += 0.02 * ev[:, 3]
        out.append((read, emap.astype(np.int64), ev))
    return table, out


@pytest.mark.gpu
def test_device_estimation_matches_golden_and_oracle_bit_for_bit():
    sa.gpu_init(0)
    for tag in ("r9p4", "zymo"):
        g, m, read, emap, events = _case(tag)
        model = sa.Model.from_tables(str(m["alphabet"]), int(m["k"]), m["trans"], m["match"], gap_y=m["gap_y"], gap_x=m["gap_x"])
        params, evs = sa.estimate_params_batch(model, [(read, emap, events)] * 3)
        for p, ev in zip(params, evs):
            assert [p[k] for k in NAMES] == list(g[tag + "_params"]), (tag, p)
            assert np.array_equal(ev[:, 0], g[tag + "_means_after"]) and np.array_equal(ev[:, 1:], events[:, 1:])
        params, evs = sa.estimate_params_batch(model, [(read, emap, events)], adjust_events=False)
        assert np.array_equal(evs[0], events) and params[0]["drift"] == g[tag + "_params"][6]
    table, strands = _synthetic_strands(64)
    model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
    params, evs = sa.estimate_params_batch(model, strands)
    for (read, emap, ev0), p, ev in zip(strands, params, evs):
        want, wev, n = port.estimate_params(table, "ACGT", 6, read, emap, ev0)
        assert n == p["n_assignments"] and [p[k] for k in NAMES] == [want[k] for k in NAMES]
        assert np.array_equal(ev, wev)
    with pytest.raises(sa.SaError, match="alphabet"):
        sa.estimate_params_batch(model, [("ACGTNACGTACGT", np.arange(13), np.ones((13, 4)))])
