"""Shared helpers for the parity tests (GPU path vs the oracle)."""
import numpy as np

import signalalign_b200 as sa
from oracle import port

TOL_P = 1e-4          # north_star: every posterior within 1e-4 absolute
TOL_P_INT = 1000      # the same in floor(p * 1e7) units


def port_model_from(alphabet, k, trans, table, kind=port.EMIT_GAUSS_MEANONLY_DESCALE, gap_y=None, **kw):
    return port.PortModel(alphabet, k, trans, table, gap_y=gap_y, kind=kind, **kw)


def oracle_align(pm, read, p):
    """read: dict(target, events, anchors, scale, shift, var[, ragged_left, ragged_right])"""
    pm.scale, pm.shift, pm.var = read["scale"], read["shift"], read["var"]
    pairs, stats = port.align(pm, read["target"], read["events"], read["anchors"], p,
                              ragged_left=read.get("ragged_left", True), ragged_right=read.get("ragged_right", True),
                              ambig=read.get("ambig"))
    return pairs, stats


def sa_read(read):
    return sa.Read(read["target"], read["events"], read["anchors"], read["scale"], read["shift"], read["var"],
                   read.get("ragged_left", True), read.get("ragged_right", True))


def port_params(p):
    return port.params(p.threshold, p.minDiagsBetweenTraceBack, p.traceBackDiagonals, p.diagonalExpansion,
                       p.splitMatrixBiggerThanThis)


def compare_pairs(gpu, ora, threshold, tol_int=TOL_P_INT):
    """gpu: int32[n,4], ora: int64[m,4] (p1e7, x, y, kmerIdx), both in the reference's list order.

    Contract (BASELINE.json north_star): integer outputs bit-exact, posteriors within 1e-4, pair set identical
    except pairs within tolerance of the threshold.  Returns max |dp| in 1e-7 units.
    """
    g = {(int(x), int(y), int(k)): int(p) for p, x, y, k in gpu}
    o = {(int(x), int(y), int(k)): int(p) for p, x, y, k in ora}
    assert len(g) == len(gpu), "duplicate (x, y, kmer) in the GPU output"
    thr_int = threshold * 1e7
    maxd = 0
    for key in set(g) | set(o):
        if key in g and key in o:
            d = abs(g[key] - o[key])
            maxd = max(maxd, d)
            assert d <= tol_int, f"posterior of pair {key} differs by {d} (1e-7 units): gpu {g[key]} oracle {o[key]}"
        else:
            p = g.get(key, o.get(key))
            assert p <= thr_int + tol_int, f"pair {key} (p={p}) present on one side only and not near the threshold"
    if len(gpu) == len(ora):
        same_order = np.array_equal(np.asarray(gpu)[:, 1:], np.asarray(ora)[:, 1:].astype(np.int32))
        assert same_order, "pair lists have the same members but a different order than the reference list"
    return maxd
