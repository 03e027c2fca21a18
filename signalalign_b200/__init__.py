"""signalalign_b200 — B200-native hot path of signalAlign (banded forward/backward posterior-match DP).

This module is a thin ctypes binding over the C-ABI of ``libsagpu.so`` (include/signalalign_b200.h).  All compute is
in the CUDA library; there is no CPU fallback and nothing here touches ``oracle/``: if the library or a CUDA device is
missing, calls raise.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsagpu.so")

EMIT_GAUSS_MEANONLY_DESCALE = 0
EMIT_HDP = 1
EMIT_GAUSS_NOISE_SCALED = 2
EMIT_GAUSS_NOISE_DESCALE = 3

c_i64p = C.POINTER(C.c_int64)
c_i32p = C.POINTER(C.c_int32)
c_f64p = C.POINTER(C.c_double)


class SaRead(C.Structure):
    _fields_ = [("target", C.c_char_p), ("targetLength", C.c_int64), ("events", c_f64p), ("nEvents", C.c_int64),
                ("anchors", c_i64p), ("nAnchors", C.c_int64), ("scale", C.c_double), ("shift", C.c_double),
                ("var", C.c_double), ("raggedLeft", C.c_int32), ("raggedRight", C.c_int32)]


class SaParams(C.Structure):
    _fields_ = [("threshold", C.c_double), ("minDiagsBetweenTraceBack", C.c_int64),
                ("traceBackDiagonals", C.c_int64), ("diagonalExpansion", C.c_int64),
                ("splitMatrixBiggerThanThis", C.c_int64)]


class SaPair(C.Structure):
    _fields_ = [("p", C.c_int32), ("x", C.c_int32), ("y", C.c_int32), ("kmerIdx", C.c_int32)]


class SaStrand(C.Structure):
    _fields_ = [("read", C.c_char_p), ("readLength", C.c_int64), ("strandEventMap", c_i64p), ("events", c_f64p),
                ("nEvents", C.c_int64)]


class SaScaling(C.Structure):
    _fields_ = [("scale", C.c_double), ("shift", C.c_double), ("var", C.c_double), ("shift_sd", C.c_double),
                ("scale_sd", C.c_double), ("var_sd", C.c_double), ("drift", C.c_double), ("nAssignments", C.c_int64)]


class SaVariantRead(C.Structure):
    _fields_ = [("pairs", C.POINTER(SaPair)), ("nPairs", C.c_int64), ("target", C.c_char_p), ("targetLength", C.c_int64),
                ("strandIsComplement", C.c_int32), ("forward", C.c_int32), ("referenceSequenceOffset", C.c_int64)]


class SaMarginals(C.Structure):
    _fields_ = [("nReads", C.c_int64), ("nVariants", C.c_int32), ("variants", C.c_char * 32), ("positionOffsets", c_i64p),
                ("positions", c_i64p), ("probs", c_f64p), ("perRead", c_f64p)]


class SaEventRead(C.Structure):
    _fields_ = [("sequence", C.c_char_p), ("sequenceLength", C.c_int64), ("eventMean", C.POINTER(C.c_float)), ("nEvents", C.c_int64),
                ("rna", C.c_int32), ("useMom", C.c_int32), ("scale", C.c_double), ("shift", C.c_double), ("var", C.c_double)]


class SaEventAlignment(C.Structure):
    _fields_ = [("nReads", C.c_int64), ("pairOffsets", c_i64p), ("kmerIdx", c_i32p), ("eventIdx", c_i32p), ("scale", c_f64p),
                ("shift", c_f64p), ("var", c_f64p), ("avgLogEmission", c_f64p), ("maxGap", c_i32p), ("spanned", c_i32p),
                ("failed", c_i32p), ("nAlignedBeforeQc", c_i32p)]


class SaResult(C.Structure):
    _fields_ = [("nReads", C.c_int64), ("pairOffsets", c_i64p), ("pairs", C.POINTER(SaPair)),
                ("transitions", c_f64p), ("likelihood", c_f64p), ("assignOffsets", c_i64p),
                ("assignKmerPos", c_i32p), ("assignEvent", c_i32p), ("totalBandCells", C.c_int64),
                ("totalPathCells", C.c_int64), ("totalEvents", C.c_int64)]


EXPORTS = [
    "sa_gpu_shutdown", "sa_model_load", "sa_model_from_tables", "sa_model_set_hdp", "sa_model_attach_nhdp",
    "sa_model_set_hdp_expected_values",
    "sa_model_scale_noise", "sa_model_scale_model", "sa_model_free", "sa_model_num_kmers", "sa_model_kmer_length",
    "sa_model_alphabet", "sa_model_get_tables", "sa_model_hdp_density", "sa_params_default", "sa_ambig_load",
    "sa_ambig_free", "sa_ambig_create", "sa_ambig_set", "sa_result_free", "sa_kmer_string", "sa_gpu_init", "sa_last_error", "sa_gpu_device_count",
    "sa_gpu_align_batch", "sa_gpu_expectations_batch", "sa_batch_prepare", "sa_batch_run", "sa_batch_fetch",
    "sa_batch_last_timing", "sa_batch_band_cells", "sa_batch_kernel_items", "sa_batch_wide_items", "sa_batch_h2d_bytes", "sa_batch_d2h_bytes", "sa_batch_free",
    "sa_result_sum_expectations", "sa_debug_logadd", "sa_model_hdp_dims", "sa_model_get_hdp", "sa_debug_band",
    "sa_debug_split_points", "sa_debug_kmer_paths", "sa_last_call_stats", "sa_set_host_threads",
    "sa_gpu_comm_unique_id", "sa_gpu_comm_init", "sa_gpu_comm_from_nccl", "sa_gpu_comm_free",
    "sa_gpu_allreduce_expectations", "sa_gpu_allgather_assignments", "sa_gpu_align_noalloc", "sa_gpu_estimate_params_batch", "sa_gpu_set_precision", "sa_gpu_get_precision",
    "sa_gpu_marginalize_variants", "sa_marginals_free", "sa_gpu_event_align_batch", "sa_event_alignment_free",
]

_lib = None


def build(verbose=False):
    """Compile libsagpu.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(HERE, "csrc")]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("SA_LIBSAGPU", LIB_PATH)   # A/B runs of differently compiled kernels (tools/ab.sh)
    if not os.path.exists(path):
        raise RuntimeError("libsagpu.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "signalalign_b200 has no CPU fallback")
    L = C.CDLL(path)
    L.sa_last_error.restype = C.c_char_p
    L.sa_model_load.restype = C.c_void_p
    L.sa_model_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    L.sa_model_from_tables.restype = C.c_void_p
    L.sa_model_from_tables.argtypes = [C.c_char_p, C.c_int32, c_f64p, c_f64p, c_f64p, c_f64p, C.c_int]
    L.sa_model_set_hdp.argtypes = [C.c_void_p, C.c_int64, c_f64p, c_i32p, c_i32p, C.c_int64, c_f64p, c_f64p]
    L.sa_model_set_hdp_expected_values.argtypes = [C.c_void_p]
    L.sa_model_attach_nhdp.argtypes = [C.c_void_p, C.c_char_p]
    L.sa_model_scale_noise.argtypes = [C.c_void_p, C.c_double, C.c_double]
    L.sa_model_scale_model.argtypes = [C.c_void_p] + [C.c_double] * 5
    L.sa_model_free.argtypes = [C.c_void_p]
    L.sa_model_num_kmers.restype = C.c_int64
    L.sa_model_num_kmers.argtypes = [C.c_void_p]
    L.sa_model_kmer_length.restype = C.c_int32
    L.sa_model_kmer_length.argtypes = [C.c_void_p]
    L.sa_model_alphabet.restype = C.c_char_p
    L.sa_model_alphabet.argtypes = [C.c_void_p]
    L.sa_model_get_tables.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p]
    L.sa_last_call_stats.restype = None
    L.sa_last_call_stats.argtypes = [C.POINTER(C.c_int64)] * 3
    L.sa_model_hdp_density.restype = C.c_double
    L.sa_model_hdp_density.argtypes = [C.c_void_p, C.c_int64, C.c_double]
    L.sa_params_default.argtypes = [C.POINTER(SaParams)]
    L.sa_ambig_load.restype = C.c_void_p
    L.sa_ambig_load.argtypes = [C.c_char_p]
    L.sa_ambig_free.argtypes = [C.c_void_p]
    L.sa_ambig_create.restype = C.c_void_p
    L.sa_ambig_set.argtypes = [C.c_void_p, C.c_char, C.c_char_p]
    L.sa_result_free.argtypes = [C.POINTER(SaResult)]
    L.sa_kmer_string.argtypes = [C.c_void_p, C.c_int64, C.c_char_p]
    L.sa_gpu_init.argtypes = [C.c_int]
    L.sa_gpu_align_batch.argtypes = [C.c_void_p, C.POINTER(SaRead), C.c_int64, C.POINTER(SaParams), C.c_void_p,
                                     C.POINTER(C.POINTER(SaResult))]
    L.sa_gpu_expectations_batch.argtypes = L.sa_gpu_align_batch.argtypes
    L.sa_batch_prepare.argtypes = [C.c_void_p, C.POINTER(SaRead), C.c_int64, C.POINTER(SaParams), C.c_void_p,
                                   C.c_int, C.POINTER(C.c_void_p)]
    L.sa_batch_run.argtypes = [C.c_void_p]
    L.sa_batch_fetch.argtypes = [C.c_void_p, C.POINTER(C.POINTER(SaResult))]
    L.sa_batch_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_float), c_i64p]
    for f in ("sa_batch_band_cells", "sa_batch_h2d_bytes", "sa_batch_d2h_bytes"):
        getattr(L, f).restype = C.c_int64
        getattr(L, f).argtypes = [C.c_void_p]
    L.sa_batch_kernel_items.argtypes = [C.c_void_p, c_i64p]
    L.sa_batch_wide_items.argtypes = [C.c_void_p, c_i64p]
    L.sa_batch_free.argtypes = [C.c_void_p]
    L.sa_gpu_align_noalloc.argtypes = [C.c_void_p, C.POINTER(SaRead), C.POINTER(SaParams), C.POINTER(SaPair), C.c_int64, c_i64p]
    L.sa_gpu_estimate_params_batch.argtypes = [C.c_void_p, C.POINTER(SaStrand), C.c_int64, C.c_int, C.POINTER(SaScaling)]
    L.sa_gpu_marginalize_variants.argtypes = [C.c_void_p, C.POINTER(SaVariantRead), C.c_int64, C.c_char_p, C.c_int,
                                              C.POINTER(C.POINTER(SaMarginals))]
    L.sa_marginals_free.restype = None
    L.sa_gpu_event_align_batch.argtypes = [C.c_void_p, C.POINTER(SaEventRead), C.c_int64, C.POINTER(C.POINTER(SaEventAlignment))]
    L.sa_event_alignment_free.restype = None
    L.sa_event_alignment_free.argtypes = [C.POINTER(SaEventAlignment)]
    L.sa_marginals_free.argtypes = [C.POINTER(SaMarginals)]
    L.sa_gpu_comm_unique_id.argtypes = [C.c_void_p]
    L.sa_gpu_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.sa_gpu_comm_from_nccl.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.sa_gpu_comm_free.argtypes = [C.c_void_p]
    L.sa_gpu_allreduce_expectations.argtypes = [C.c_void_p, c_f64p, C.c_int64]
    L.sa_gpu_allgather_assignments.argtypes = [C.c_void_p, c_i32p, c_i32p, C.c_int64, C.POINTER(c_i32p), C.POINTER(c_i32p), c_i64p]
    L.sa_result_sum_expectations.argtypes = [C.POINTER(SaResult), c_f64p]
    L.sa_debug_logadd.argtypes = [c_f64p, c_f64p, c_f64p, C.c_int64]
    L.sa_model_hdp_dims.argtypes = [C.c_void_p, c_i64p]
    L.sa_model_get_hdp.argtypes = [C.c_void_p, c_f64p, c_i32p, c_i32p, c_f64p, c_f64p]
    L.sa_debug_band.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, c_i64p]
    L.sa_debug_split_points.restype = C.c_int64
    L.sa_debug_split_points.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, c_i64p,
                                        C.c_int64]
    L.sa_debug_kmer_paths.restype = C.c_int64
    L.sa_debug_kmer_paths.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, c_i32p, C.c_int64]
    _lib = L
    return L


class SaError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise SaError(lib().sa_last_error().decode(errors="replace"))


def gpu_init(device=0):
    _check(lib().sa_gpu_init(device))


def gpu_shutdown():
    _check(lib().sa_gpu_shutdown())


def default_params(**kw):
    """pairwiseAlignmentBandingParameters_construct defaults (reference impl/pairwiseAligner.c:2022-2037)."""
    p = SaParams()
    lib().sa_params_default(C.byref(p))
    names = {"threshold": "threshold", "min_diags": "minDiagsBetweenTraceBack", "trace_back": "traceBackDiagonals",
             "expansion": "diagonalExpansion", "split_bigger": "splitMatrixBiggerThanThis"}
    for k, v in kw.items():
        setattr(p, names[k], v)
    return p


class Model:
    """Host mirror of the reference `StateMachine` for the path (reference inc/stateMachine.h:85-191)."""

    def __init__(self, handle):
        if not handle:
            raise SaError(lib().sa_last_error().decode(errors="replace"))
        self.h = handle
        self.K = lib().sa_model_num_kmers(self.h)
        self.k = lib().sa_model_kmer_length(self.h)
        self.alphabet = lib().sa_model_alphabet(self.h).decode()

    @classmethod
    def load(cls, model_file, nhdp_file=None, kind=EMIT_GAUSS_MEANONLY_DESCALE):
        return cls(lib().sa_model_load(model_file.encode(), nhdp_file.encode() if nhdp_file else None, kind))

    @classmethod
    def from_tables(cls, alphabet, k, log_transitions, match, gap_y=None, gap_x=None,
                    kind=EMIT_GAUSS_MEANONLY_DESCALE):
        t = np.ascontiguousarray(log_transitions, np.float64)
        m = np.ascontiguousarray(match, np.float64)
        gy = None if gap_y is None else np.ascontiguousarray(gap_y, np.float64)
        gx = None if gap_x is None else np.ascontiguousarray(gap_x, np.float64)
        f = lambda a: None if a is None else a.ctypes.data_as(c_f64p)
        return cls(lib().sa_model_from_tables(alphabet.encode(), k, f(t), f(m), f(gy), f(gx), kind))

    def set_hdp(self, grid, kmer_to_row, post_pred, slopes, own_row=None):
        grid = np.ascontiguousarray(grid, np.float64)
        k2r = np.ascontiguousarray(kmer_to_row, np.int32)
        own = None if own_row is None else np.ascontiguousarray(own_row, np.int32)
        pp = np.ascontiguousarray(post_pred, np.float64)
        sl = np.ascontiguousarray(slopes, np.float64)
        _check(lib().sa_model_set_hdp(self.h, len(grid), grid.ctypes.data_as(c_f64p), k2r.ctypes.data_as(c_i32p),
                                      None if own is None else own.ctypes.data_as(c_i32p), pp.shape[0], pp.ctypes.data_as(c_f64p), sl.ctypes.data_as(c_f64p)))

    def set_hdp_expected_values(self):
        _check(lib().sa_model_set_hdp_expected_values(self.h))

    def attach_nhdp(self, nhdp_file):
        _check(lib().sa_model_attach_nhdp(self.h, nhdp_file.encode()))

    def scale_noise(self, scale_sd, var_sd):
        _check(lib().sa_model_scale_noise(self.h, scale_sd, var_sd))

    def scale_model(self, scale, shift, var, scale_sd, var_sd):
        _check(lib().sa_model_scale_model(self.h, scale, shift, var, scale_sd, var_sd))

    def tables(self):
        t = np.zeros(9)
        m = np.zeros((self.K, 5))
        y = np.zeros((self.K, 5))
        lib().sa_model_get_tables(self.h, t.ctypes.data_as(c_f64p), m.ctypes.data_as(c_f64p),
                                  y.ctypes.data_as(c_f64p))
        return t, m, y

    def hdp_density(self, kmer_idx, x):
        return lib().sa_model_hdp_density(self.h, kmer_idx, x)

    def hdp_tables(self):
        """(grid[G], kmer_to_row[K], post_pred[rows, G], slopes[rows, G], own_row[K]) as loaded from the .nhdp"""
        dims = (C.c_int64 * 2)()
        lib().sa_model_hdp_dims(self.h, dims)
        G, rows = dims[0], dims[1]
        grid = np.zeros(G)
        k2r = np.zeros(self.K, np.int32)
        own = np.zeros(self.K, np.int32)
        pp = np.zeros((rows, G))
        sl = np.zeros((rows, G))
        lib().sa_model_get_hdp(self.h, grid.ctypes.data_as(c_f64p), k2r.ctypes.data_as(c_i32p),
                               own.ctypes.data_as(c_i32p), pp.ctypes.data_as(c_f64p), sl.ctypes.data_as(c_f64p))
        return grid, k2r, pp, sl, own

    def kmer_paths(self, kmer, ambig=None, cap=4096):
        """hdCell_construct2 path enumeration of one k-window (host-side; no device needed)"""
        out = np.zeros(cap, np.int32)
        n = lib().sa_debug_kmer_paths(self.h, ambig.h if ambig else None, kmer.encode(), out.ctypes.data_as(c_i32p), cap)
        if n < 0:
            raise SaError("kmer %r contains a character outside the alphabet %s" % (kmer, self.alphabet))
        return out[:n].copy()

    def kmer_string(self, idx):
        buf = C.create_string_buffer(self.k + 1)
        lib().sa_kmer_string(self.h, idx, buf)
        return buf.value.decode()

    def kmer_index(self, kmer):
        a = len(self.alphabet)
        idx = 0
        for ch in kmer:
            idx = idx * a + self.alphabet.index(ch)
        return idx

    def __del__(self):
        try:
            if self.h:
                lib().sa_model_free(self.h)
                self.h = None
        except Exception:
            pass


class Read:
    """Arguments of one getAlignedPairsUsingAnchors call (reference inc/pairwiseAligner.h:406-414)."""

    def __init__(self, target, events, anchors, scale=1.0, shift=0.0, var=1.0, ragged_left=True, ragged_right=True):
        self.target = target.encode() if isinstance(target, str) else bytes(target)
        self.events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
        self.anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
        self.scale, self.shift, self.var = float(scale), float(shift), float(var)
        self.ragged_left, self.ragged_right = bool(ragged_left), bool(ragged_right)
        self._c = None

    def cstruct(self):
        """the sa_read_t over this read's buffers (built once: the arrays are owned by the object and never re-bound)"""
        if self._c is None:
            self._c = self._build()
        return self._c

    def _build(self):
        return SaRead(self.target, len(self.target), self.events.ctypes.data_as(c_f64p), self.events.shape[0],
                      self.anchors.ctypes.data_as(c_i64p), self.anchors.shape[0], self.scale, self.shift, self.var,
                      int(self.ragged_left), int(self.ragged_right))


def _read_array(reads):
    arr = (SaRead * max(1, len(reads)))()
    for i, r in enumerate(reads):
        arr[i] = r.cstruct()
    return arr


class Ambig:
    def __init__(self, path=None, table=None):
        """path: a `code<TAB>bases` file (None = the built-in table); table: a dict code -> bases instead of a file"""
        if table is not None:
            self.h = lib().sa_ambig_create()
            for code, bases in table.items():
                _check(lib().sa_ambig_set(self.h, code.encode(), bases.encode()))
        else:
            self.h = lib().sa_ambig_load(path.encode() if path else None)
        if not self.h:
            raise SaError(lib().sa_last_error().decode(errors="replace"))

    def __del__(self):
        try:
            lib().sa_ambig_free(self.h)
        except Exception:
            pass


class Result:
    """Owns an sa_result_t; the arrays are zero-copy numpy views that stay valid while this object lives."""

    def __init__(self, ptr):
        self._ptr = ptr
        r = ptr.contents
        n = r.nReads
        self.n_reads = n
        self.band_cells, self.path_cells, self.events = r.totalBandCells, r.totalPathCells, r.totalEvents
        self.pair_offsets = np.ctypeslib.as_array(r.pairOffsets, shape=(n + 1,))
        total = int(self.pair_offsets[-1])
        if total and r.pairs:
            self.pairs = np.ctypeslib.as_array(C.cast(r.pairs, c_i32p), shape=(total, 4))
        else:
            self.pairs = np.zeros((0, 4), np.int32)
        self.transitions = self.likelihood = None
        self.assign_offsets = self.assign_kmer_pos = self.assign_event = None
        if r.transitions:
            self.transitions = np.ctypeslib.as_array(r.transitions, shape=(n, 9)) if n else np.zeros((0, 9))
            self.likelihood = np.ctypeslib.as_array(r.likelihood, shape=(n,)) if n else np.zeros(0)
        if r.assignOffsets:
            self.assign_offsets = np.ctypeslib.as_array(r.assignOffsets, shape=(n + 1,))
            ta = int(self.assign_offsets[-1])
            self.assign_kmer_pos = np.ctypeslib.as_array(r.assignKmerPos, shape=(max(ta, 1),))[:ta]
            self.assign_event = np.ctypeslib.as_array(r.assignEvent, shape=(max(ta, 1),))[:ta]

    def read_pairs(self, i):
        """pairs of read i as int32 [n, 4] = (floor(p*1e7), x, y, kmerIdx) in the reference's list order"""
        return self.pairs[self.pair_offsets[i]:self.pair_offsets[i + 1]]

    def __del__(self):
        try:
            if self._ptr is not None:
                self.pairs = self.pair_offsets = self.transitions = self.likelihood = None
                self.assign_offsets = self.assign_kmer_pos = self.assign_event = None
                lib().sa_result_free(self._ptr)
                self._ptr = None
        except Exception:
            pass


def align_batch(model, reads, params=None, ambig=None):
    """Batch of getAlignedPairsUsingAnchors calls (host buffers in, host result out)."""
    params = params or default_params()
    arr = _read_array(reads)
    out = C.POINTER(SaResult)()
    _check(lib().sa_gpu_align_batch(model.h, arr, len(reads), C.byref(params), ambig.h if ambig else None,
                                    C.byref(out)))
    return Result(out)


def align_noalloc(model, read, params=None, pair_cap=None):
    """sa_gpu_align_noalloc: one read, no device allocation (static device scratch); int32 [n, 4] pairs like read_pairs"""
    params = params or default_params()
    cap = int(pair_cap if pair_cap is not None else 3 * read.events.shape[0] + 64)
    out = np.zeros((max(cap, 1), 4), np.int32)
    n = C.c_int64(0)
    rs = read.cstruct()
    _check(lib().sa_gpu_align_noalloc(model.h, C.byref(rs), C.byref(params), C.cast(out.ctypes.data, C.POINTER(SaPair)), cap,
                                      C.byref(n)))
    return out[:n.value].copy()


def estimate_params_batch(model, strands, adjust_events=True):
    """signalUtils_estimateNanoporeParams for a batch: strands = [(read string, event map int64[len(read)], events [n, 4])].

    Returns (list of dicts scale / shift / var / shift_sd / scale_sd / var_sd / drift / n_assignments, list of the events
    arrays after the drift adjustment)."""
    keep, arr = [], (SaStrand * max(1, len(strands)))()
    for i, (read, emap, events) in enumerate(strands):
        rd = read.encode() if isinstance(read, str) else bytes(read)
        em = np.ascontiguousarray(emap, np.int64)
        ev = np.array(events, np.float64, order="C", copy=True).reshape(-1, 4)
        keep.append((rd, em, ev))
        arr[i] = SaStrand(rd, len(rd), em.ctypes.data_as(c_i64p), ev.ctypes.data_as(c_f64p), ev.shape[0])
    out = (SaScaling * max(1, len(strands)))()
    _check(lib().sa_gpu_estimate_params_batch(model.h, arr, len(strands), int(adjust_events), out))
    names = ["scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift"]
    params = [dict({k: getattr(out[i], k) for k in names}, n_assignments=out[i].nAssignments) for i in range(len(strands))]
    return params, [k[2] for k in keep]


def marginalize_variants(model, reads, variants, round_like_tsv=True):
    """Variant calls of aligned pairs and their per-position marginals (sa_gpu_marginalize_variants).

    reads = [(pairs int32 [n, 4] as Result.read_pairs returns them, target string with X at the ambiguous positions,
    strand_is_complement, forward, reference_sequence_offset)].  Returns (sorted variants, list of
    (positions int64 [m], probs float64 [m, len(variants)], per_read float64 [len(variants)]))."""
    keep, arr = [], (SaVariantRead * max(1, len(reads)))()
    for i, (pairs, target, comp, fwd, off) in enumerate(reads):
        pr = np.ascontiguousarray(pairs, np.int32).reshape(-1, 4)
        tg = target.encode() if isinstance(target, str) else bytes(target)
        keep.append((pr, tg))
        arr[i] = SaVariantRead(pr.ctypes.data_as(C.POINTER(SaPair)), pr.shape[0], tg, len(tg), int(bool(comp)), int(bool(fwd)), int(off))
    out = C.POINTER(SaMarginals)()
    vb = variants.encode() if isinstance(variants, str) else bytes(variants)
    _check(lib().sa_gpu_marginalize_variants(model.h, arr, len(reads), vb, int(bool(round_like_tsv)), C.byref(out)))
    try:
        r = out.contents
        V = r.nVariants
        offs = np.ctypeslib.as_array(r.positionOffsets, (len(reads) + 1,)).copy()
        tot = int(offs[-1])
        pos = np.ctypeslib.as_array(r.positions, (max(1, tot),))[:tot].copy()
        probs = np.ctypeslib.as_array(r.probs, (max(1, tot) * V,))[:tot * V].reshape(tot, V).copy()
        per = np.ctypeslib.as_array(r.perRead, (max(1, len(reads)) * V,)).reshape(-1, V)[:len(reads)].copy()
        res = [(pos[offs[i]:offs[i + 1]], probs[offs[i]:offs[i + 1]], per[i]) for i in range(len(reads))]
        return r.variants.decode(), res
    finally:
        lib().sa_marginals_free(out)


def event_align_batch(model, reads):
    """Adaptive banded event -> kmer alignment (sa_gpu_event_align_batch).

    reads = [dict(sequence=str, event_mean=float32 array, rna=False, scalings=None | (scale, shift, var))]; scalings None =
    the reference's method of moments.  Returns a list of dicts: pairs int32 [n, 2] = (kmer, event), scale, shift, var,
    avg_log_emission, max_gap, spanned, failed, n_aligned_before_qc."""
    keep, arr = [], (SaEventRead * max(1, len(reads)))()
    for i, r in enumerate(reads):
        sq = r["sequence"].encode() if isinstance(r["sequence"], str) else bytes(r["sequence"])
        ev = np.ascontiguousarray(r["event_mean"], np.float32)
        keep.append((sq, ev))
        sc = r.get("scalings")
        arr[i] = SaEventRead(sq, len(sq), ev.ctypes.data_as(C.POINTER(C.c_float)), ev.shape[0], int(bool(r.get("rna"))),
                             int(sc is None), *(sc if sc is not None else (1.0, 0.0, 1.0)))
    out = C.POINTER(SaEventAlignment)()
    _check(lib().sa_gpu_event_align_batch(model.h, arr, len(reads), C.byref(out)))
    try:
        r = out.contents
        n = len(reads)
        offs = np.ctypeslib.as_array(r.pairOffsets, (n + 1,)).copy()
        tot = int(offs[-1])
        km = np.ctypeslib.as_array(r.kmerIdx, (max(1, tot),))[:tot].copy()
        ei = np.ctypeslib.as_array(r.eventIdx, (max(1, tot),))[:tot].copy()
        col = lambda p, t=np.float64: np.ctypeslib.as_array(p, (max(1, n),))[:n].copy()
        sc, sh, va, av = col(r.scale), col(r.shift), col(r.var), col(r.avgLogEmission)
        mg, sp, fl, na = col(r.maxGap), col(r.spanned), col(r.failed), col(r.nAlignedBeforeQc)
        return [dict(pairs=np.stack([km[offs[i]:offs[i + 1]], ei[offs[i]:offs[i + 1]]], 1), scale=float(sc[i]), shift=float(sh[i]),
                     var=float(va[i]), avg_log_emission=float(av[i]), max_gap=int(mg[i]), spanned=bool(sp[i]), failed=bool(fl[i]),
                     n_aligned_before_qc=int(na[i])) for i in range(n)]
    finally:
        lib().sa_event_alignment_free(out)


def expectations_batch(model, reads, params=None, ambig=None):
    """Batch of getExpectationsUsingAnchors calls; per-read sums WITHOUT pseudocounts."""
    params = params or default_params()
    arr = _read_array(reads)
    out = C.POINTER(SaResult)()
    _check(lib().sa_gpu_expectations_batch(model.h, arr, len(reads), C.byref(params), ambig.h if ambig else None,
                                           C.byref(out)))
    return Result(out)


class Comm:
    """NCCL communicator behind the C-ABI (sa_gpu_comm_*): the all-reduce of the expectation tables and the gather of the
    HDP assignment lists.  `unique_id()` on rank 0, carry the 128 bytes to the other ranks, then Comm(id, n, rank)."""

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        _check(lib().sa_gpu_comm_unique_id(buf))
        return buf.raw

    def __init__(self, unique_id, n_ranks, rank):
        self.h = C.c_void_p()
        self.n_ranks, self.rank = n_ranks, rank
        _check(lib().sa_gpu_comm_init(C.c_char_p(unique_id), n_ranks, rank, C.byref(self.h)))

    def allreduce_expectations(self, table):
        """in-place sum over ranks of a float64 table (same length on every rank); returns it"""
        t = np.ascontiguousarray(table, np.float64)
        _check(lib().sa_gpu_allreduce_expectations(self.h, t.ctypes.data_as(c_f64p), t.size))
        return t

    def allgather_assignments(self, kmer_pos, event):
        k = np.ascontiguousarray(kmer_pos, np.int32)
        e = np.ascontiguousarray(event, np.int32)
        ak, ae = c_i32p(), c_i32p()
        counts = np.zeros(self.n_ranks, np.int64)
        _check(lib().sa_gpu_allgather_assignments(self.h, k.ctypes.data_as(c_i32p), e.ctypes.data_as(c_i32p), k.size,
                                                  C.byref(ak), C.byref(ae), counts.ctypes.data_as(c_i64p)))
        total = int(counts.sum())
        outk = np.ctypeslib.as_array(ak, shape=(max(total, 1),))[:total].copy()
        oute = np.ctypeslib.as_array(ae, shape=(max(total, 1),))[:total].copy()
        libc = C.CDLL(None)
        libc.free.argtypes = [C.c_void_p]
        libc.free(ak)
        libc.free(ae)
        return outk, oute, counts

    def close(self):
        if self.h:
            lib().sa_gpu_comm_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def set_precision(name):
    """'fp64' (default: the reference's double arithmetic) or 'fp32' (wide band kernel with float state relative to
    per-diagonal offsets, Gaussian alignment runs; posteriors within 1e-4)"""
    _check(lib().sa_gpu_set_precision({"fp64": 0, "fp32": 1}[name]))


def get_precision():
    return ["fp64", "fp32"][lib().sa_gpu_get_precision()]


def set_host_threads(n):
    """host threads for the packing passes of this process (launchers like torchrun export OMP_NUM_THREADS=1)"""
    lib().sa_set_host_threads(int(n))


def last_call_stats():
    """(H2D bytes, D2H bytes, kernel launches) of the last align_batch call of this process"""
    a, b, c = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    lib().sa_last_call_stats(C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def debug_logadd(x, y):
    x = np.ascontiguousarray(x, np.float64)
    y = np.ascontiguousarray(y, np.float64)
    out = np.zeros_like(x)
    _check(lib().sa_debug_logadd(x.ctypes.data_as(c_f64p), y.ctypes.data_as(c_f64p), out.ctypes.data_as(c_f64p),
                                 x.size))
    return out


def debug_band(anchors, lX, lY, expansion):
    """band_construct on the host (reference impl/pairwiseAligner.c:195-246): int64[lX+lY+1, 3] = xay, xmyL, xmyR"""
    a = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    out = np.zeros((lX + lY + 1, 3), np.int64)
    lib().sa_debug_band(a.ctypes.data_as(c_i64p), a.shape[0], lX, lY, expansion, out.ctypes.data_as(c_i64p))
    return out


def debug_split_points(anchors, lX, lY, split_bigger, ragged_left, ragged_right):
    """getSplitPoints on the host (reference impl/pairwiseAligner.c:1910-1937): int64[n, 4] = x1, y1, x2, y2"""
    a = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    cap = a.shape[0] + 4
    out = np.zeros((cap, 4), np.int64)
    n = lib().sa_debug_split_points(a.ctypes.data_as(c_i64p), a.shape[0], lX, lY, split_bigger, int(ragged_left),
                                    int(ragged_right), out.ctypes.data_as(c_i64p), cap)
    return out[:n].copy()


class Batch:
    """Staged form: prepare (pack + H2D), run (kernels only; repeatable), fetch (D2H)."""

    def __init__(self, model, reads, params=None, ambig=None, expectations=False):
        self.model, self.reads, self.ambig = model, reads, ambig  # keep host buffers alive
        self.params = params or default_params()
        arr = _read_array(reads)
        self.h = C.c_void_p()
        _check(lib().sa_batch_prepare(model.h, arr, len(reads), C.byref(self.params), ambig.h if ambig else None,
                                      int(expectations), C.byref(self.h)))

    def run(self):
        _check(lib().sa_batch_run(self.h))

    def fetch(self):
        out = C.POINTER(SaResult)()
        _check(lib().sa_batch_fetch(self.h, C.byref(out)))
        return Result(out)

    def timing(self):
        ms = (C.c_float * 3)()
        n = C.c_int64(0)
        lib().sa_batch_last_timing(self.h, ms, C.byref(n))
        return [ms[0], ms[1], ms[2]], n.value

    @property
    def band_cells(self):
        return lib().sa_batch_band_cells(self.h)

    @property
    def kernel_items(self):
        """work items by kernel: (general, band 32/64 slots, band 128/256 slots, wide)"""
        o = (C.c_int64 * 4)()
        lib().sa_batch_kernel_items(self.h, o)
        return tuple(o)

    @property
    def wide_items(self):
        """work items of the wide band kernel by window class (256, 512, 1024, 2048, 4096 slots)"""
        o = (C.c_int64 * 5)()
        lib().sa_batch_wide_items(self.h, o)
        return tuple(o)

    @property
    def h2d_bytes(self):
        return lib().sa_batch_h2d_bytes(self.h)

    @property
    def d2h_bytes(self):
        return lib().sa_batch_d2h_bytes(self.h)

    def close(self):
        if self.h:
            lib().sa_batch_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
