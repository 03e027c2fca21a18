"""ctypes bindings for oracle/libsa_oracle.so (the CPU restatement, oracle/sa_oracle.c) — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsa_oracle.so")

EMIT_GAUSS_MEANONLY_DESCALE = 0
EMIT_HDP = 1
EMIT_GAUSS_NOISE_SCALED = 2
EMIT_GAUSS_NOISE_DESCALE = 3

c_i64p = C.POINTER(C.c_int64)
c_f64p = C.POINTER(C.c_double)
c_i32p = C.POINTER(C.c_int32)


class SaoModel(C.Structure):
    _fields_ = [("kind", C.c_int32), ("alphabetSize", C.c_int32), ("kmerLength", C.c_int32), ("pad", C.c_int32),
                ("K", C.c_int64), ("alphabet", C.c_char * 64), ("trans", C.c_double * 9),
                ("matchTable", c_f64p), ("gapYTable", c_f64p), ("gapXTable", c_f64p),
                ("scale", C.c_double), ("shift", C.c_double), ("var", C.c_double),
                ("gridLength", C.c_int64), ("grid", c_f64p), ("kmerToRow", c_i32p),
                ("postPred", c_f64p), ("slopes", c_f64p)]


class SaoParams(C.Structure):
    _fields_ = [("threshold", C.c_double), ("minDiagsBetweenTraceBack", C.c_int64),
                ("traceBackDiagonals", C.c_int64), ("diagonalExpansion", C.c_int64),
                ("splitMatrixBiggerThanThis", C.c_int64)]


class SaoAmbig(C.Structure):
    _fields_ = [("opts", (C.c_char * 16) * 256)]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.sao_logAdd.restype = C.c_double
        L.sao_logAdd.argtypes = [C.c_double, C.c_double]
        L.sao_kmer_id.restype = C.c_int64
        L.sao_kmer_id.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_int64]
        L.sao_emission.restype = C.c_double
        L.sao_emission.argtypes = [C.POINTER(SaoModel), C.c_int64, C.c_double, C.c_double, C.c_int]
        L.sao_band.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, c_i64p]
        L.sao_split_points.restype = C.c_int64
        L.sao_split_points.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                       c_i64p, C.c_int64]
        L.sao_paths.restype = C.c_int64
        L.sao_paths.argtypes = [C.c_char_p, C.POINTER(SaoModel), C.POINTER(SaoAmbig), c_i32p, C.c_int64]
        L.sao_ambig_default.argtypes = [C.POINTER(SaoAmbig)]
        L.sao_ambig_from_file.argtypes = [C.POINTER(SaoAmbig), C.c_char_p]
        L.sao_align.restype = C.c_int64
        L.sao_align.argtypes = [C.POINTER(SaoModel), C.c_char_p, C.c_int64, c_f64p, C.c_int64, c_i64p, C.c_int64,
                                C.POINTER(SaoParams), C.c_int, C.c_int, C.POINTER(SaoAmbig), c_i64p, C.c_int64,
                                c_f64p]
        L.sao_expectations.restype = C.c_int64
        L.sao_expectations.argtypes = [C.POINTER(SaoModel), C.c_char_p, C.c_int64, c_f64p, C.c_int64, c_i64p,
                                       C.c_int64, C.POINTER(SaoParams), C.c_int, C.c_int, C.POINTER(SaoAmbig),
                                       c_f64p, c_f64p, c_i64p, c_i64p, C.c_int64]
        L.sao_align_noband.restype = C.c_int64
        L.sao_align_noband.argtypes = [C.POINTER(SaoModel), C.c_char_p, C.c_int64, c_f64p, C.c_int64,
                                       C.POINTER(SaoParams), C.c_int, C.c_int, C.POINTER(SaoAmbig), c_i64p,
                                       C.c_int64]
        _lib = L
    return _lib


def _i64(a):
    return a.ctypes.data_as(c_i64p)


def _f64(a):
    return a.ctypes.data_as(c_f64p)


def params(threshold=0.01, min_diags=1000, trace_back=40, expansion=20, split_bigger=3000 * 3000):
    """Library defaults of pairwiseAlignmentBandingParameters_construct (pairwiseAligner.c:2022-2037)."""
    return SaoParams(threshold, min_diags, trace_back, expansion, split_bigger)


class PortModel:
    """Flat model for the oracle port. Tables are numpy arrays kept alive by this object."""

    def __init__(self, alphabet, k, trans_log, match, gap_y=None, gap_x=None, kind=EMIT_GAUSS_MEANONLY_DESCALE,
                 scale=1.0, shift=0.0, var=1.0, grid=None, kmer_to_row=None, post_pred=None, slopes=None):
        self.alphabet = "".join(sorted(alphabet))
        self.k = k
        self.K = len(alphabet) ** k
        self.match = np.ascontiguousarray(match, np.float64).reshape(self.K, 5).copy()
        if gap_y is None:
            gap_y = self.match.copy()
            gap_y[:, 1] *= 1.75  # EXTRA_EVENT_NOISE_MULTIPLIER, stateMachine.c:1528-1532
        self.gap_y = np.ascontiguousarray(gap_y, np.float64).reshape(self.K, 5).copy()
        if gap_x is None:
            gap_x = np.full(self.K, -2.3025850929940455)
        self.gap_x = np.ascontiguousarray(gap_x, np.float64)
        self.trans = np.ascontiguousarray(trans_log, np.float64)
        self.kind = kind
        self.scale, self.shift, self.var = scale, shift, var
        self.grid = None if grid is None else np.ascontiguousarray(grid, np.float64)
        self.kmer_to_row = None if kmer_to_row is None else np.ascontiguousarray(kmer_to_row, np.int32)
        self.post_pred = None if post_pred is None else np.ascontiguousarray(post_pred, np.float64)
        self.slopes = None if slopes is None else np.ascontiguousarray(slopes, np.float64)

    def cstruct(self):
        m = SaoModel()
        m.kind = self.kind
        m.alphabetSize = len(self.alphabet)
        m.kmerLength = self.k
        m.K = self.K
        m.alphabet = self.alphabet.encode()
        for i in range(9):
            m.trans[i] = self.trans[i]
        m.matchTable = _f64(self.match)
        m.gapYTable = _f64(self.gap_y)
        m.gapXTable = _f64(self.gap_x)
        m.scale, m.shift, m.var = self.scale, self.shift, self.var
        if self.grid is not None:
            m.gridLength = len(self.grid)
            m.grid = _f64(self.grid)
            m.kmerToRow = self.kmer_to_row.ctypes.data_as(c_i32p)
            m.postPred = _f64(self.post_pred)
            m.slopes = _f64(self.slopes)
        return m

    def kmer_string(self, idx):
        a = len(self.alphabet)
        s = []
        for _ in range(self.k):
            s.append(self.alphabet[idx % a])
            idx //= a
        return "".join(reversed(s))


def load_model_file(path, kind=EMIT_GAUSS_MEANONLY_DESCALE):
    """Parse a signalAlign .model file (format: stateMachine.c:1445-1452; transitions mapping :1202-1258)."""
    with open(path) as f:
        l0 = f.readline().split()
        l1 = f.readline().split()
        l2 = f.readline().split()
    n_states, a_size, alphabet, k = int(l0[0]), int(l0[1]), l0[2], int(l0[3])
    assert n_states == 3 and len(alphabet) == a_size
    t = [float(v) for v in l1]
    assert len(t) == 10
    with np.errstate(divide="ignore"):
        lg = np.log(np.array(t[:9]))
    # order: MATCH_CONTINUE, MATCH_FROM_GAP_X, MATCH_FROM_GAP_Y, GAP_OPEN_X, GAP_OPEN_Y, GAP_EXTEND_X, GAP_EXTEND_Y,
    #        GAP_SWITCH_TO_X (never loaded: stays LOG_ZERO), GAP_SWITCH_TO_Y (file token 7)
    trans = np.array([lg[0], lg[3], lg[6], lg[1], lg[2], lg[4], lg[8], -np.inf, lg[7]])
    table = np.array([float(v) for v in l2]).reshape(-1, 5)
    assert table.shape[0] == a_size ** k
    return PortModel(alphabet, k, trans, table, kind=kind)


def make_ambig(path=None):
    a = SaoAmbig()
    if path:
        if lib().sao_ambig_from_file(C.byref(a), path.encode()) != 0:
            raise IOError(path)
    else:
        lib().sao_ambig_default(C.byref(a))
    return a


def align(model, target, events, anchors, p=None, ragged_left=True, ragged_right=True, ambig=None):
    """Returns (pairs[n,4] int64 (p1e7, x, y, kmerIdx), stats[4])."""
    p = p or params()
    events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    lY = events.shape[0]
    lX = max(0, len(target) - (model.k - 1))
    m = model.cstruct()
    amb = ambig or make_ambig()
    cap = 64 * (lX + lY) + 1024
    stats = np.zeros(4)
    while True:
        out = np.zeros((cap, 4), np.int64)
        n = lib().sao_align(C.byref(m), target.encode(), lX, _f64(events), lY, _i64(anchors), anchors.shape[0],
                            C.byref(p), int(ragged_left), int(ragged_right), C.byref(amb), _i64(out), cap,
                            _f64(stats))
        if n < 0:
            raise ValueError("kmer contains a character outside the alphabet")
        if n <= cap:
            return out[:n].copy(), stats
        cap = n


def align_noband(model, target, events, p=None, ragged_left=False, ragged_right=False, ambig=None):
    p = p or params()
    events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
    lY = events.shape[0]
    lX = len(target) - (model.k - 1)
    m = model.cstruct()
    amb = ambig or make_ambig()
    cap = 64 * (lX + lY) + 1024
    out = np.zeros((cap, 4), np.int64)
    n = lib().sao_align_noband(C.byref(m), target.encode(), lX, _f64(events), lY, C.byref(p), int(ragged_left),
                               int(ragged_right), C.byref(amb), _i64(out), cap)
    return out[:n].copy()


def expectations(model, target, events, anchors, p=None, ragged_left=True, ragged_right=True, ambig=None,
                 pseudo_t=0.001):
    """Returns (transitions9 incl. pseudocount, likelihood, assignKmerOff, assignEvent)."""
    p = p or params()
    events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    lY = events.shape[0]
    lX = len(target) - (model.k - 1)
    m = model.cstruct()
    amb = ambig or make_ambig()
    cap = 8 * (lX + lY) + 1024
    while True:
        tr = np.full(9, pseudo_t)
        lik = np.zeros(1)
        ak = np.zeros(cap, np.int64)
        ae = np.zeros(cap, np.int64)
        n = lib().sao_expectations(C.byref(m), target.encode(), lX, _f64(events), lY, _i64(anchors), anchors.shape[0],
                                   C.byref(p), int(ragged_left), int(ragged_right), C.byref(amb), _f64(tr), _f64(lik),
                                   _i64(ak), _i64(ae), cap)
        if n <= cap:
            break
        cap = n
    if n < 0:
        raise ValueError("kmer contains a character outside the alphabet")
    return tr, float(lik[0]), ak[:n].copy(), ae[:n].copy()


def band(anchors, lX, lY, expansion):
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    out = np.zeros((lX + lY + 1, 3), np.int64)
    lib().sao_band(_i64(anchors), anchors.shape[0], lX, lY, expansion, _i64(out))
    return out


def split_points(anchors, lX, lY, split_bigger, ragged_left, ragged_right):
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    cap = anchors.shape[0] + 4
    out = np.zeros((cap, 4), np.int64)
    n = lib().sao_split_points(_i64(anchors), anchors.shape[0], lX, lY, split_bigger, int(ragged_left),
                               int(ragged_right), _i64(out), cap)
    return out[:n].copy()


def log_add(x, y):
    return lib().sao_logAdd(x, y)


def kmer_id(kmer, alphabet):
    return lib().sao_kmer_id(kmer.encode(), alphabet.encode(), len(alphabet), len(kmer))


def emission(model, kmer_idx, mean, noise=1.0, match=True):
    m = model.cstruct()
    return lib().sao_emission(C.byref(m), kmer_idx, mean, noise, 1 if match else 0)


def estimate_params(match_table, alphabet, k, read, event_map, events, adjust_events=True):
    """signalUtils_estimateNanoporeParams (reference impl/signalMachineUtils.c:186-225) on one strand.

    Returns (params dict, events after the drift adjustment, number of assignments); alphabet must be sorted."""
    t = np.ascontiguousarray(match_table, np.float64)
    em = np.ascontiguousarray(event_map, np.int64)
    ev = np.array(events, np.float64, order="C", copy=True).reshape(-1, 4)
    out = np.zeros(7)
    L = lib()
    L.sao_estimate_params.restype = C.c_int64
    L.sao_estimate_params.argtypes = [c_f64p, C.c_char_p, C.c_int64, C.c_int64, C.c_char_p, C.c_int64, c_i64p, c_f64p,
                                      C.c_int64, C.c_int, c_f64p]
    rd = read.encode() if isinstance(read, str) else bytes(read)
    n = L.sao_estimate_params(_f64(t), alphabet.encode(), len(alphabet), k, rd, len(rd), _i64(em), _f64(ev), ev.shape[0],
                              int(adjust_events), _f64(out))
    names = ["scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift"]
    return dict(zip(names, out)), ev, int(n)


# ---------------------------------------------------------------------------------------------------- SURVEY 8f N3
# variant calls of one strand of one read and their per-position marginals (plain Python / numpy: small cases only)
_RC = {"A": "T", "C": "G", "G": "C", "T": "A", "a": "t", "c": "g", "g": "c", "t": "a", "M": "K", "K": "M", "R": "Y", "Y": "R",
       "B": "V", "V": "B", "D": "H", "H": "D", "m": "k", "k": "m", "r": "y", "y": "r", "b": "v", "v": "b", "d": "h", "h": "d"}


def tsv_round6(p):
    """the value a `%f` field carries (the reference's pipeline passes every posterior through the TSV text)"""
    return float("%f" % p)


def variant_rows(pairs, target, k, kmer_strings, strand_complement=False, forward=True, event_offset=0, ref_offset=0, rna=False):
    """writePosteriorProbsVC (reference impl/signalMachine.c:161-232): for every aligned pair (p*1e7, x, y, kmer) whose
    reference kmer holds the ambiguity character X (AMBIG_BASE, inc/pairwiseAligner.h:32; path_findDegeneratePositions
    impl/pairwiseAligner.c:577-592) one row per X: (event index, reference position, called base, p).
    pairs: int array [n, 4]; kmer_strings[i]: the path kmer of pair i.  Rows in the writer's order."""
    ref_len = len(target)
    ref_len_kmers = ref_len - k
    same = (not strand_complement and forward) or (strand_complement and not forward)      # :57, :65, :82
    rows = []
    for (p, x, y, _), path_kmer in zip(pairs, kmer_strings):
        k_i = target[x:x + k]
        ref_kmer = k_i if same else "".join(_RC.get(c, c) for c in reversed(k_i))            # makeReferenceKmer :64-70
        q_pos = [i for i, c in enumerate(ref_kmer) if c == "X"]
        if not q_pos:
            continue
        x_adj = x + ref_offset if same else ref_len_kmers - (x + (ref_len - ref_offset))     # adjustReferenceCoordinate :54-62
        for q in q_pos:
            qq = q if same else (k - 1) - q                                                  # adjustQueryPosition :81-87
            rows.append((int(y) + event_offset, x_adj + q, path_kmer[qq], int(p) / 1e7))
    return rows


def marginalize_variants(rows, variants, round_like_tsv=True):
    """MarginalizeVariants.get_data for one (strand, mapping) group (reference src/signalalign/variantCaller.py:37-89):
    per reference position the posterior mass of every called base, normalised over the bases; per read the mean over the
    positions.  Returns (sorted positions, probs [n, len(variants)] in sorted(variants) order, per-read means)."""
    vs = sorted(variants)
    mass = {}
    for _, pos, base, p in rows:
        if round_like_tsv:
            p = tsv_round6(p)
        mass.setdefault(pos, {}).setdefault(base, []).append(p)
    positions = sorted(mass)
    probs = np.zeros((len(positions), len(vs)))
    for i, pos in enumerate(positions):
        sums = {b: sum(v) for b, v in mass[pos].items()}          # Python's left-to-right sum, as the reference's sum()
        total = 0.0
        for b in sums:
            total += sums[b]
        for b, s in sums.items():
            probs[i, vs.index(b)] = s / total                     # (a base outside `variants` raises, as the reference does)
    per_read = probs.sum(0) / max(1, len(positions))
    return np.array(positions, dtype=np.int64), probs, per_read


# ---------------------------------------------------------------------------------------------------- SURVEY 8f N4
def mom_scalings(match_table, kmer_ids, event_means):
    """estimate_scalings_using_mom (reference impl/eventAligner.c:792-842): shift and scale by the method of moments (the
    events are C floats, the sums sequential doubles); var = 1"""
    ev = np.asarray(event_means, np.float32).astype(np.float64)
    lm = np.asarray(match_table, np.float64).reshape(-1, 5)[np.asarray(kmer_ids), 0]
    s = 0.0
    for v in ev:
        s += v
    ks = ksq = 0.0
    for v in lm:
        ks += v
        ksq += v * v              # pow(levelMean, 2.0)
    shift = s / len(ev) - ks / len(lm)
    sq = 0.0
    for v in ev:
        sq += (v - shift) * (v - shift)
    scale = (sq / len(ev)) / (ksq / len(lm))
    return scale, shift, 1.0


def event_align(match_table, kmer_ids, event_means, scale, shift, var):
    """adaptive_banded_simple_event_align2 (reference impl/eventAligner.c:899-1235): Suzuki-Kasahara adaptive band of 100
    cells over (event, kmer), Viterbi with step / stay / skip, trim states, backtrack and the reference's QC.  One numpy
    pass per band.  Returns (aligned (kmer, event) pairs int64 [n, 2] — empty when QC fails —, dict of QC figures)."""
    tbl = np.asarray(match_table, np.float64).reshape(-1, 5)
    kid = np.asarray(kmer_ids, np.int64)
    ev = np.asarray(event_means, np.float32).astype(np.float64)
    n_kmers, n_events = len(kid), len(ev)
    mu, sigma = tbl[kid, 0], tbl[kid, 1]
    c1 = -0.91893853320467267 - np.log(sigma)            # log_inv_sqrt_2pi - l_sigma (stateMachine.c:296-306)
    log_inv_var = np.log(1 / var)

    def emission(e_idx, k_idx):                          # ..._MeanOnly (stateMachine.c:557-605), operation order kept
        m, s = mu[k_idx], sigma[k_idx]
        x = (ev[e_idx] + var * m - scale * m - shift) / var
        a = (x - m) / s
        return log_inv_var + (c1[k_idx] + (-0.5 * a * a))

    FROM_D, FROM_U, FROM_L = 0, 1, 2
    bw, half = 100, 50
    events_per_kmer = n_events / n_kmers
    p_stay = 1 - (1 / (events_per_kmer + 1))
    lp_skip = np.log(1e-10)
    lp_stay = np.log(p_stay)
    lp_step = np.log(1.0 - np.exp(lp_skip) - np.exp(lp_stay))
    lp_trim = np.log(0.01)
    n_bands = (n_events + 1) + (n_kmers + 1)
    bands = np.full((n_bands, bw), -np.inf)
    trace = np.zeros((n_bands, bw), np.uint8)
    ll_e = np.zeros(n_bands, np.int64)
    ll_k = np.zeros(n_bands, np.int64)
    ll_e[0], ll_k[0] = half - 1, -1 - half
    ll_e[1], ll_k[1] = ll_e[0] + 1, ll_k[0]
    bands[0, -1 - ll_k[0]] = 0.0
    o = ll_e[1] - 0
    bands[1, o] = lp_trim
    trace[1, o] = FROM_U
    off = np.arange(bw)
    f32 = lambda a: np.asarray(a, np.float64).astype(np.float32).astype(np.float64)
    for b in range(2, n_bands):
        ll, ur = bands[b - 1, 0], bands[b - 1, bw - 1]
        if ll == -np.inf and ur == -np.inf:
            right = b % 2 == 1
        else:
            right = ll < ur
        ll_e[b], ll_k[b] = (ll_e[b - 1], ll_k[b - 1] + 1) if right else (ll_e[b - 1] + 1, ll_k[b - 1])
        t_off = -1 - ll_k[b]
        if 0 <= t_off < bw:
            e_idx = ll_e[b] - t_off
            if 0 <= e_idx < n_events:
                bands[b, t_off] = lp_trim * (e_idx + 1)
                trace[b, t_off] = FROM_U
            else:
                bands[b, t_off] = -np.inf
        lo = max(0 - ll_k[b], ll_e[b] - (n_events - 1), 0)
        hi = min(n_kmers - ll_k[b], ll_e[b] - (-1), bw)
        if hi <= lo:
            continue
        o = off[lo:hi]
        e_idx, k_idx = ll_e[b] - o, ll_k[b] + o
        o_up = ll_e[b - 1] - (e_idx - 1)
        o_left = (k_idx - 1) - ll_k[b - 1]
        o_diag = (k_idx - 1) - ll_k[b - 2]

        def fetch(row, oo):
            ok = (oo >= 0) & (oo < bw)
            return np.where(ok, f32(bands[row, np.clip(oo, 0, bw - 1)]), -np.inf)       # float up / left / diag
        up, left, diag = fetch(b - 1, o_up), fetch(b - 1, o_left), fetch(b - 2, o_diag)
        lp = emission(e_idx, k_idx)
        with np.errstate(invalid="ignore"):
            s_d, s_u, s_l = f32(diag + lp_step + lp), f32(up + lp_stay + lp), f32(left + lp_skip)
        mx = s_d.copy()
        fr = np.full(len(o), FROM_D, np.uint8)
        mx = np.where(s_u > mx, s_u, mx)
        fr = np.where(mx == s_u, FROM_U, fr)
        mx = np.where(s_l > mx, s_l, mx)
        fr = np.where(mx == s_l, FROM_L, fr)
        bands[b, lo:hi] = mx
        trace[b, lo:hi] = fr
    # backtrack
    k_cur = n_kmers - 1
    best, e_cur = np.float32(-np.inf), 0
    for e_idx in range(n_events):
        b = (e_idx + 1) + (k_cur + 1)
        oo = ll_e[b] - e_idx
        if 0 <= oo < bw:
            s = np.float32(bands[b, oo] + (n_events - e_idx) * lp_trim)
            if s > best:
                best, e_cur = s, e_idx
    out, sum_em, n_al, gap, max_gap = [], 0.0, 0.0, 0, 0
    while k_cur >= 0 and e_cur >= 0:
        out.append((k_cur, e_cur))
        sum_em += float(emission(np.array([e_cur]), np.array([k_cur]))[0])
        n_al += 1
        b = (e_cur + 1) + (k_cur + 1)
        fr = trace[b, ll_e[b] - e_cur]
        if fr == FROM_D:
            k_cur -= 1; e_cur -= 1; gap = 0
        elif fr == FROM_U:
            e_cur -= 1; gap = 0
        else:
            k_cur -= 1; gap += 1; max_gap = max(gap, max_gap)
    out.reverse()
    avg = sum_em / n_al
    spanned = out[0][0] == 0 and out[-1][0] == n_kmers - 1
    failed = avg < -5.2 or not spanned or max_gap > 50 or events_per_kmer > 5.0
    qc = dict(avg_log_emission=avg, spanned=bool(spanned), max_gap=int(max_gap), failed=bool(failed), n_aligned=len(out))
    return (np.zeros((0, 2), np.int64) if failed else np.array(out, np.int64)), qc
