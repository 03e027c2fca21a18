"""Shared helpers for the parity tests (GPU path / oracle port vs the golden fixtures and vs each other)."""
import os

import numpy as np

import signalalign_b200 as sa
from oracle import port

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT_DIR = os.path.dirname(HERE)
GOLDEN = os.path.join(HERE, "golden")

TOL_P = 1e-4          # north_star: every posterior within 1e-4 absolute
TOL_P_INT = 1000      # the same in floor(p * 1e7) units
TOL_LIK_REL = 1e-5    # north_star: total log-likelihood within 1e-5 relative

_cache = {}


def golden(name):
    """tests/golden/<name>.npz as a dict (generated from the compiled reference by tests/golden/make_golden.py)"""
    if name not in _cache:
        with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
            _cache[name] = {k: z[k] for k in z.files}
    return _cache[name]


def params_from(vec):
    """golden params vector = threshold, minDiags, traceBack, expansion, splitBigger"""
    return dict(threshold=float(vec[0]), min_diags=int(vec[1]), trace_back=int(vec[2]), expansion=int(vec[3]),
                split_bigger=int(vec[4]))


def scale_noise_tables(match, gap_y, scale_sd, var_sd):
    """emissions_signal_scaleNoise (reference impl/stateMachine.c:721-741) on copies of the K x 5 tables"""
    m, y = match.copy(), gap_y.copy()
    m[:, 2] = m[:, 2] * scale_sd
    m[:, 4] = m[:, 4] * var_sd
    m[:, 3] = np.sqrt(np.power(m[:, 2], 3.0) / m[:, 4])
    y[:, 2] = y[:, 2] * scale_sd
    y[:, 4] = y[:, 4] * var_sd
    y[:, 3] = np.sqrt(np.power(y[:, 2], 3.0) / m[:, 4])
    return m, y


def scale_model_tables(match, gap_y, scale, shift, var, scale_sd, var_sd):
    """emissions_signal_scaleModel (reference impl/stateMachine.c:743-779)"""
    m, y = match.copy(), gap_y.copy()
    m[:, 0] = m[:, 0] * scale + shift
    m[:, 1] = m[:, 1] * var
    y[:, 0] = y[:, 0] * scale + shift
    y[:, 1] = y[:, 1] * var
    return scale_noise_tables(m, y, scale_sd, var_sd)


def golden_models(model_name, kind, tparams=None, hdp=False):
    """(PortModel, sa.Model-or-None) for a golden case, prepared the way the reference test prepared its StateMachine.

    kind 0: production MeanOnly de-scaling; 2: scaled tables + noise (loadScaledStateMachine3); 3: de-scaling + noise with
    scaleNoise (loadDescaledStateMachine3); 1 / hdp: HDP density.  The sa.Model is only built when libsagpu.so loads.
    """
    g = golden(model_name)
    alphabet, k = str(g["alphabet"]), int(g["k"])
    match, gap_y, gap_x, trans = g["match"], g["gap_y"], g["gap_x"], g["trans"]
    if kind == port.EMIT_GAUSS_NOISE_SCALED:
        match, gap_y = scale_model_tables(match, gap_y, tparams[0], tparams[1], tparams[2], tparams[4], tparams[5])
    elif kind == port.EMIT_GAUSS_NOISE_DESCALE:
        match, gap_y = scale_noise_tables(match, gap_y, tparams[4], tparams[5])
    kw = {}
    if hdp:
        h = golden("hdp_single_level_fixed")
        kw = dict(grid=h["grid"], kmer_to_row=h["kmer_to_row"], post_pred=h["post_pred"], slopes=h["slopes"])
        kind = port.EMIT_HDP
    pm = port.PortModel(alphabet, k, trans, match, gap_y=gap_y, gap_x=gap_x, kind=kind, **kw)
    try:
        sm = sa.Model.from_tables(alphabet, k, trans, match, gap_y=gap_y, gap_x=gap_x, kind=kind)
        if hdp:
            sm.set_hdp(kw["grid"], kw["kmer_to_row"], kw["post_pred"], kw["slopes"], own_row=h["own_row"])
    except Exception:
        sm = None
    return pm, sm


def port_model_from(alphabet, k, trans, table, kind=port.EMIT_GAUSS_MEANONLY_DESCALE, gap_y=None, **kw):
    return port.PortModel(alphabet, k, trans, table, gap_y=gap_y, kind=kind, **kw)


def oracle_align(pm, read, p):
    """read: dict(target, events, anchors, scale, shift, var[, ragged_left, ragged_right])"""
    pm.scale, pm.shift, pm.var = read["scale"], read["shift"], read["var"]
    pairs, stats = port.align(pm, read["target"], read["events"], read["anchors"], p,
                              ragged_left=read.get("ragged_left", True), ragged_right=read.get("ragged_right", True),
                              ambig=read.get("ambig"))
    return pairs, stats


def oracle_expectations(pm, read, p):
    pm.scale, pm.shift, pm.var = read["scale"], read["shift"], read["var"]
    return port.expectations(pm, read["target"], read["events"], read["anchors"], p,
                             ragged_left=read.get("ragged_left", True), ragged_right=read.get("ragged_right", True),
                             ambig=read.get("ambig"), pseudo_t=0.0)


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


def compare_expectations(got_t, got_l, want_t, want_l, pseudo=0.0):
    """transitions: relative 1e-6 (absolute 1e-9 floor for the structurally-zero entries); likelihood: 1e-5 relative"""
    want = np.asarray(want_t, np.float64) - pseudo
    got = np.asarray(got_t, np.float64)
    # (a read whose events fall where the HDP density is 0 makes the reference itself return NaN sums: same places)
    assert np.allclose(got, want, rtol=1e-6, atol=1e-9, equal_nan=True), (got, want)
    if not np.isfinite(want_l):
        assert (np.isnan(want_l) and np.isnan(got_l)) or got_l == want_l, (got_l, want_l)
    else:
        assert abs(got_l - want_l) <= TOL_LIK_REL * abs(want_l), (got_l, want_l)


def write_model_file(path, alphabet, k, trans_log, table):
    """A .model file in the reference's 3-line text format (impl/stateMachine.c:1445-1452) from raw tables.

    trans_log is in StateMachine3 order; the file wants row-major M,X,Y x M,X,Y probabilities with token 5 unused and
    token 7 = GAP_SWITCH_TO_Y (stateMachine3_loadTransitionsFromFile, impl/stateMachine.c:1202-1258)."""
    t = np.exp(np.asarray(trans_log, np.float64))
    row = [t[0], t[3], t[4], t[1], t[5], 0.0, t[2], t[8], t[6]]
    with open(path, "w") as f:
        f.write("3\t%d\t%s\t%d\n" % (len(alphabet), alphabet, k))
        f.write("\t".join(repr(float(v)) for v in row) + "\t0.0\n")
        f.write("\t".join(repr(float(v)) for v in np.asarray(table).reshape(-1)) + "\n")


def write_nhdp_file(path, alphabet, k, grid, post_pred_by_dp, slopes_by_dp, parents, data_dp_ids):
    """A .nhdp in the reference's text format (impl/nanopore_hdp.c:1045-1086 + impl/hdp.c:2919-3050).

    post_pred_by_dp / slopes_by_dp: dict dp id -> array[G] for the DPs that carry a distribution; parents: list of
    parent id or -1 per DP; data_dp_ids: DP id of every data point (decides which DPs are `observed`)."""
    n = len(parents)
    G = len(grid)
    fmt = lambda a: "\t".join("%.17g" % float(v) for v in a)
    with open(path, "w") as f:
        f.write("%d\n%s\n%d\n" % (len(alphabet), alphabet, k))
        f.write("1\n1\n0\n%d\n" % n)                       # splines_finalized, has_data, sample_gamma, num_dps
        f.write(fmt(np.full(len(data_dp_ids), 70.0)) + "\n")
        f.write("\t".join(str(int(v)) for v in data_dp_ids) + "\n")
        f.write("%.17g\t%.17g\t%.17g\t%.17g\n" % (70.0, 1.0, 2.0, 10.0))   # mu nu alpha beta
        f.write("%.17g\t%.17g\t%d\n" % (grid[0], grid[-1], G))
        f.write(fmt([1.0, 1.0]) + "\n")                     # gamma per depth
        for pid in parents:
            f.write(("-" if pid < 0 else str(int(pid))) + "\t0\n")
        for table in (post_pred_by_dp, slopes_by_dp):
            for i in range(n):
                f.write((fmt(table[i]) if i in table else "") + "\n")
        # no factor lines: the density query never reads the factor tree, and deserialize_hdp (impl/hdp.c:3262-3318)
        # takes every further line, blank ones included, for a factor


def write_golden_hdp_files(dirname):
    """The ACEGOT 6-mer .model / .nhdp pair of BASELINE config 2 rebuilt from the golden tables (model_r73_acegot,
    hdp_single_level_fixed) in the reference's own text formats, for tests that run the REFERENCE's loaders where
    /root/reference does not exist.  Single level: every kmer DP hangs off the root, DP id K."""
    h = golden("hdp_single_level_fixed")
    t = golden("model_r73_acegot")
    own = h["own_row"]
    K = len(own)
    seen = np.nonzero(own >= 0)[0]
    root_row = int(h["kmer_to_row"][own < 0][0])
    pp = {int(k): h["post_pred"][own[k]] for k in seen}
    sl = {int(k): h["slopes"][own[k]] for k in seen}
    pp[K], sl[K] = h["post_pred"][root_row], h["slopes"][root_row]
    model_file = os.path.join(dirname, "acegot.model")
    nhdp_file = os.path.join(dirname, "single_level_fixed.nhdp")
    write_model_file(model_file, str(t["alphabet"]), int(t["k"]), t["trans"], t["match"])
    write_nhdp_file(nhdp_file, str(t["alphabet"]), int(t["k"]), h["grid"], pp, sl, [K] * K + [-1], seen)
    return model_file, nhdp_file
