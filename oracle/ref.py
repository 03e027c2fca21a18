"""ctypes bindings for oracle/_ref/libsaref.so — TEST INFRASTRUCTURE ONLY.

libsaref.so is the UNMODIFIED reference hot path (compiled by oracle/Makefile from /root/reference) plus
oracle/ref_harness.c.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
arm may import this module; nothing under signalalign_b200/ does.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
# SA_REF_LIB: load another build of the same harness instead — tests/test_gpu_shim.py points it at libsaref_gpu.so (the
# reference with its two entry points bound to libsagpu) in a process of its own
LIB_PATH = os.environ.get("SA_REF_LIB") or os.path.join(REF_DIR, "libsaref.so")
GPU_LIB_PATH = os.path.join(REF_DIR, "libsaref_gpu.so")

_lib = None

c_i64p = C.POINTER(C.c_int64)
c_f64p = C.POINTER(C.c_double)


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        L.ref_sm_load.restype = C.c_void_p
        L.ref_sm_load.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        L.ref_sm_set_scaling.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.ref_sm_scale_model.argtypes = [C.c_void_p] + [C.c_double] * 5
        L.ref_sm_scale_noise.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.ref_sm_hdp_expected_values.argtypes = [C.c_void_p]
        L.ref_sm_info.argtypes = [C.c_void_p, c_i64p, C.c_char_p, c_f64p, c_f64p]
        L.ref_sm_tables.argtypes = [C.c_void_p, c_f64p, c_f64p, c_f64p]
        L.ref_sm_emission.restype = C.c_double
        L.ref_sm_emission.argtypes = [C.c_void_p, C.c_char_p, C.c_double, C.c_double, C.c_int]
        L.ref_hdp_density.restype = C.c_double
        L.ref_hdp_density.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.ref_kmer_id.restype = C.c_int64
        L.ref_kmer_id.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_int64]
        L.ref_logAdd.restype = C.c_double
        L.ref_logAdd.argtypes = [C.c_double, C.c_double]
        L.ref_band.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, c_i64p]
        L.ref_split_points.restype = C.c_int64
        L.ref_split_points.argtypes = [c_i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                       c_i64p, C.c_int64]
        L.ref_align.restype = C.c_int64
        L.ref_align.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, c_f64p, C.c_int64, c_i64p, C.c_int64,
                                C.c_double, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                C.c_char_p, c_i64p, C.c_char_p, C.c_int64, c_f64p]
        L.ref_align_noband.restype = C.c_int64
        L.ref_align_noband.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, c_f64p, C.c_int64, C.c_double,
                                       C.c_int, C.c_int, c_i64p, C.c_int64]
        L.ref_expectations.restype = C.c_int64
        L.ref_expectations.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, c_f64p, C.c_int64, c_i64p, C.c_int64,
                                       C.c_double, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                       C.c_char_p, C.c_double, C.c_double, c_f64p, c_f64p, c_i64p, c_i64p,
                                       C.c_int64, c_f64p]
        L.ref_npread_load.restype = C.c_void_p
        L.ref_npread_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int64]
        L.ref_npread_info.argtypes = [C.c_void_p, c_i64p, c_f64p]
        for f in ("ref_npread_twoD", "ref_npread_template_read"):
            getattr(L, f).restype = C.c_char_p
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_npread_template_events.restype = c_f64p
        L.ref_npread_template_events.argtypes = [C.c_void_p]
        for f in ("ref_npread_template_event_map", "ref_npread_template_strand_event_map"):
            getattr(L, f).restype = c_i64p
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_npread_descale.argtypes = [C.c_void_p]
        L.ref_estimate_params_template.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_lastz_anchors.restype = C.c_int64
        L.ref_lastz_anchors.argtypes = [C.c_char_p, C.c_char_p, c_i64p, C.c_char_p, c_i64p, C.c_int64]
        L.ref_cigar_anchors.restype = C.c_int64
        L.ref_cigar_anchors.argtypes = [C.c_char_p, C.c_int64, c_i64p, c_i64p, c_i64p, C.c_int64]
        _lib = L
    return _lib


def _i64(a):
    return a.ctypes.data_as(c_i64p)


def _f64(a):
    return a.ctypes.data_as(c_f64p)


class RefStateMachine:
    """A reference `StateMachine*` (inc/stateMachine.h:85) built by the reference's own loaders."""

    def __init__(self, model_file, kind=0, nhdp_file=None):
        self.h = lib().ref_sm_load(model_file.encode(), kind, (nhdp_file or "").encode())
        if not self.h:
            raise RuntimeError("reference failed to load " + model_file)
        info = np.zeros(6, np.int64)
        alpha = C.create_string_buffer(64)
        sc = np.zeros(3)
        tr = np.zeros(9)
        lib().ref_sm_info(self.h, _i64(info), alpha, _f64(sc), _f64(tr))
        self.type, self.K, self.alphabet_size, self.k = (int(info[0]), int(info[1]), int(info[2]), int(info[3]))
        self.alphabet = alpha.value.decode()
        self.transitions = tr

    def set_scaling(self, scale, shift, var):
        lib().ref_sm_set_scaling(self.h, scale, shift, var)

    def scale_model(self, scale, shift, var, scale_sd, var_sd):
        lib().ref_sm_scale_model(self.h, scale, shift, var, scale_sd, var_sd)

    def scale_noise(self, scale_sd, var_sd):
        lib().ref_sm_scale_noise(self.h, scale_sd, var_sd)

    def hdp_expected_values(self):
        lib().ref_sm_hdp_expected_values(self.h)

    def scaling(self):
        info = np.zeros(6, np.int64)
        alpha = C.create_string_buffer(64)
        sc = np.zeros(3)
        tr = np.zeros(9)
        lib().ref_sm_info(self.h, _i64(info), alpha, _f64(sc), _f64(tr))
        return sc

    def tables(self):
        m = np.zeros((self.K, 5))
        y = np.zeros((self.K, 5))
        x = np.zeros(self.K)
        lib().ref_sm_tables(self.h, _f64(m), _f64(y), _f64(x))
        return m, y, x

    def emission(self, kmer, mean, noise=1.0, match=True):
        return lib().ref_sm_emission(self.h, kmer.encode(), mean, noise, 1 if match else 0)

    def hdp_density(self, kmer, x):
        return lib().ref_hdp_density(self.h, kmer.encode(), x)

    def align(self, target, events, anchors, threshold=0.01, min_diags=0, trace_back=0, expansion=-1,
              split_bigger=0, ragged_left=True, ragged_right=True, ambig_file=None):
        """getAlignedPairsUsingAnchors. Returns (pairs[n,3] int64 (p1e7,x,y), kmers list[str], seconds)."""
        events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
        anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
        lY = events.shape[0]
        lX = len(target) - (self.k - 1) if len(target) > 0 else 0
        cap = max(16, 64 * (lY + lX) + 1024)
        while True:
            out = np.zeros((cap, 3), np.int64)
            km = C.create_string_buffer(cap * (self.k + 1))
            sec = C.c_double(0)
            n = lib().ref_align(self.h, target.encode(), lX, _f64(events), lY, _i64(anchors), anchors.shape[0],
                                threshold, min_diags, trace_back, expansion, split_bigger,
                                int(ragged_left), int(ragged_right), (ambig_file or "").encode(),
                                _i64(out), km, cap, C.byref(sec))
            if n <= cap:
                break
            cap = n
        raw = km.raw
        kmers = [raw[i * (self.k + 1):i * (self.k + 1) + self.k].decode() for i in range(n)]
        return out[:n].copy(), kmers, sec.value

    def align_noband(self, target, events, threshold=0.01, ragged_left=False, ragged_right=False):
        events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
        lY = events.shape[0]
        lX = len(target) - (self.k - 1)
        cap = 64 * (lX + lY) + 1024
        out = np.zeros((cap, 3), np.int64)
        n = lib().ref_align_noband(self.h, target.encode(), lX, _f64(events), lY, threshold,
                                   int(ragged_left), int(ragged_right), _i64(out), cap)
        return out[:n].copy()

    def expectations(self, target, events, anchors, threshold=0.01, min_diags=0, trace_back=0, expansion=-1,
                     split_bigger=0, ragged_left=True, ragged_right=True, ambig_file=None,
                     pseudo_t=0.001, pseudo_e=0.001):
        events = np.ascontiguousarray(events, np.float64).reshape(-1, 4)
        anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
        lY = events.shape[0]
        lX = len(target) - (self.k - 1)
        cap = 8 * (lX + lY) + 1024
        while True:
            tr = np.zeros(9)
            lik = C.c_double(0)
            ak = np.zeros(cap, np.int64)
            ae = np.zeros(cap, np.int64)
            sec = C.c_double(0)
            n = lib().ref_expectations(self.h, target.encode(), lX, _f64(events), lY, _i64(anchors), anchors.shape[0],
                                       threshold, min_diags, trace_back, expansion, split_bigger,
                                       int(ragged_left), int(ragged_right), (ambig_file or "").encode(),
                                       pseudo_t, pseudo_e, _f64(tr), C.byref(lik), _i64(ak), _i64(ae), cap,
                                       C.byref(sec))
            if n <= cap:
                break
            cap = n
        return tr, lik.value, ak[:min(n, cap)].copy(), ae[:min(n, cap)].copy(), sec.value


class RefNpRead:
    """A reference `NanoporeRead*` (inc/nanopore.h:23) loaded by nanopore_loadNanoporeReadFromFile."""

    def __init__(self, path, alphabet, alphabet_size=None):
        self.h = lib().ref_npread_load(path.encode(), alphabet.encode(), alphabet_size or len(alphabet))
        self._refresh()

    def _refresh(self):
        lens = np.zeros(5, np.int64)
        tp = np.zeros(7)
        lib().ref_npread_info(self.h, _i64(lens), _f64(tp))
        self.read_length, self.template_read_length, _, self.n_template_events, _ = [int(v) for v in lens]
        self.template_params = dict(zip(["scale", "shift", "var", "shift_sd", "scale_sd", "var_sd", "drift"], tp))
        td = lib().ref_npread_twoD(self.h)
        self.twoD_read = td.decode(errors="replace") if td and self.read_length > 0 else ""
        tr = lib().ref_npread_template_read(self.h)
        self.template_read = tr.decode(errors="replace") if tr else ""

    def template_events(self):
        p = lib().ref_npread_template_events(self.h)
        return np.ctypeslib.as_array(p, shape=(self.n_template_events, 4)).copy()

    def template_event_map(self):
        p = lib().ref_npread_template_event_map(self.h)
        return np.ctypeslib.as_array(p, shape=(self.read_length,)).copy()

    def template_strand_event_map(self):
        p = lib().ref_npread_template_strand_event_map(self.h)
        return np.ctypeslib.as_array(p, shape=(self.template_read_length,)).copy()

    def descale(self):
        lib().ref_npread_descale(self.h)
        self._refresh()

    def estimate_params(self, sm):
        lib().ref_estimate_params_template(sm.h, self.h)
        self._refresh()


def lastz_anchors(ref_seq, read_seq, event_map):
    em = np.ascontiguousarray(event_map, np.int64)
    cap = len(read_seq) + 16
    out = np.zeros((cap, 2), np.int64)
    n = lib().ref_lastz_anchors(ref_seq.encode(), read_seq.encode(), _i64(em), REF_DIR.encode(), _i64(out), cap)
    if n < 0:
        raise RuntimeError("lastz anchors failed")
    return out[:n].copy()


def band(anchors, lX, lY, expansion):
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    out = np.zeros((lX + lY + 1, 3), np.int64)
    lib().ref_band(_i64(anchors), anchors.shape[0], lX, lY, expansion, _i64(out))
    return out


def split_points(anchors, lX, lY, split_bigger, ragged_left, ragged_right):
    anchors = np.ascontiguousarray(anchors, np.int64).reshape(-1, 2)
    cap = anchors.shape[0] + 4
    out = np.zeros((cap, 4), np.int64)
    n = lib().ref_split_points(_i64(anchors), anchors.shape[0], lX, lY, split_bigger, int(ragged_left),
                               int(ragged_right), _i64(out), cap)
    return out[:n].copy()


def log_add(x, y):
    return lib().ref_logAdd(x, y)


def kmer_id(kmer, alphabet):
    return lib().ref_kmer_id(kmer.encode(), alphabet.encode(), len(alphabet), len(kmer))
