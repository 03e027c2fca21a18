"""Synthetic workloads of the shape BASELINE.json names (adapted — see DESIGN.md §Measurement).

Everything is generated with numpy from a fixed seed so the CPU oracle and the GPU path see the same bytes.
No reference file is read here: the 6-mer Gaussian table is synthetic (level means/sds drawn to resemble an ONT
R9.4 table), which keeps the bench runnable on a box without /root/reference.
"""
import numpy as np

LOG_ZERO = -np.inf

# Guide-anchor patterns (make_read's block_mean / trim).  "contract" is SURVEY.md §8(d)'s config-3 generator: 12 % of the
# positions of the true diagonal start a 1-5 nt indel, i.e. match runs of geometric length 1 / 0.12, each trimmed by 14
# on both sides (constraintDiagonalTrim, reference impl/pairwiseAligner.c:1624-1658, default :2029) — sparse anchors,
# bands hundreds to thousands of cells wide.  "long_runs_trim14": match runs of ~80 (a guide alignment whose mismatches
# stay inside M operations) with the same trim; "dense": round 1's pattern, every kmer that owns an event is an anchor.
ANCHOR_PATTERNS = {
    "contract": dict(trim=14, block_mean=1.0 / 0.12),
    "long_runs_trim14": dict(trim=14, block_mean=80),
    "dense": dict(trim=0, block_mean=80),
}


def nanopore_transitions():
    """3x3 transition probabilities in `.model` line-1 order -> the 9 log-transitions in StateMachine3 order.

    Values follow the shape of models/testModelR9.4_450bps.nucleotide.6mer.template.model (match-heavy, short gaps);
    token 7 lands in GAP_SWITCH_TO_Y and GAP_SWITCH_TO_X stays -inf exactly like
    stateMachine3_loadTransitionsFromFile (reference impl/stateMachine.c:1202-1258).
    """
    t = np.array([0.79, 0.1965, 0.0133, 0.8035, 0.1965, 0.0, 0.9867, 1e-9, 0.0133])
    with np.errstate(divide="ignore"):
        lg = np.log(t)
    return np.array([lg[0], lg[3], lg[6], lg[1], lg[2], lg[4], lg[8], LOG_ZERO, lg[7]])


def gaussian_table(alphabet="ACGT", k=6, seed=7):
    """K x 5 table: level_mean, level_sd, noise_mean, noise_sd, noise_lambda."""
    rng = np.random.default_rng(seed)
    K = len(alphabet) ** k
    t = np.zeros((K, 5))
    t[:, 0] = rng.uniform(60.0, 125.0, K)
    t[:, 1] = rng.uniform(1.2, 3.0, K)
    t[:, 2] = rng.uniform(0.8, 2.0, K)
    t[:, 4] = rng.uniform(5.0, 40.0, K)
    t[:, 3] = np.sqrt(t[:, 2] ** 3 / t[:, 4])
    return t


def kmer_indices(seq_codes, k, A):
    """rolling base-A index of every k-window of an integer-coded sequence"""
    n = len(seq_codes) - k + 1
    idx = np.zeros(n, np.int64)
    for j in range(k):
        idx = idx * A + seq_codes[j:j + n]
    return idx


def make_read(rng, ref_codes, table, alphabet="ACGT", k=6, mean_len=6000, sigma_len=0.25, min_len=2000,
              max_len=20000, block_mean=80, trim=14, n_probs=(0.10, 0.60, 0.22, 0.08)):
    """One synthetic read against a slice of the reference.

    events: walk the kmers of the slice, each emits n in {0,1,2,3} events (n_probs); mean = scale*mu + shift +
    var*sd*N(0,1).  Guide anchors: the true (kmer, first event) pairs, cut into blocks of geometric length
    `block_mean` with 1-5 kmers dropped between blocks (the guide alignment's indels) and `trim` pairs removed at
    both ends of each block (constraintDiagonalTrim, reference impl/pairwiseAligner.c:1624-1658), kept strictly
    increasing in both coordinates like filterToRemoveOverlap does.
    """
    A = len(alphabet)
    L = int(np.clip(np.round(rng.lognormal(np.log(mean_len), sigma_len)), min_len, max_len))
    L = min(L, len(ref_codes) - 1)
    start = int(rng.integers(0, len(ref_codes) - L))
    codes = ref_codes[start:start + L]
    kidx = kmer_indices(codes, k, A)
    lX = len(kidx)
    n_ev = rng.choice(4, size=lX, p=n_probs)
    if n_ev.sum() == 0:
        n_ev[0] = 1
    first = np.concatenate([[0], np.cumsum(n_ev)[:-1]])
    lY = int(n_ev.sum())
    owner = np.repeat(np.arange(lX), n_ev)
    scale, shift, var = rng.uniform(0.95, 1.05), rng.uniform(-3, 3), rng.uniform(0.9, 1.3)
    mu, sd = table[kidx[owner], 0], table[kidx[owner], 1]
    events = np.zeros((lY, 4))
    events[:, 0] = scale * mu + shift + var * sd * rng.standard_normal(lY)
    events[:, 1] = 1.0
    events[:, 2] = 0.002
    events[:, 3] = np.arange(lY) * 0.002
    # anchors
    keep = np.zeros(lX, bool)
    x = 0
    while x < lX:
        blen = int(rng.geometric(1.0 / block_mean))
        lo, hi = x + trim, min(lX, x + blen) - trim
        if hi > lo:
            keep[lo:hi] = True
        x += blen + int(rng.integers(1, 6))
    keep &= n_ev > 0
    ax = np.nonzero(keep)[0]
    ay = first[ax]
    anchors = np.stack([ax, ay], 1).astype(np.int64) if len(ax) else np.zeros((0, 2), np.int64)
    target = "".join(alphabet[c] for c in codes)
    return dict(target=target, events=events, anchors=anchors, scale=scale, shift=shift, var=var)


def make_batch(n_reads, seed=20261104, ref_len=5_000_000, alphabet="ACGT", k=6, table=None, **kw):
    rng = np.random.default_rng(seed)
    if table is None:
        table = gaussian_table(alphabet, k)
    ref_len = int(min(ref_len, max(50_000, n_reads * 20_000)))
    ref_codes = rng.integers(0, len(alphabet), ref_len).astype(np.int64)
    return [make_read(rng, ref_codes, table, alphabet, k, **kw) for _ in range(n_reads)], table
