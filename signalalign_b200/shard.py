"""Multi-GPU plumbing of the path: reads shard across ranks, nothing is exchanged on the alignment
path; the training pass (getExpectationsUsingAnchors over many reads) ends with ONE all-reduce of the per-GPU
expectation table.

One process per GPU (torchrun); `torch.distributed` is only the transport: NCCL over NVLink on the GPU box, gloo in
the CPU tests.  What is reduced mirrors what signalAlign's trainer sums over per-read expectation files
(reference src/signalalign/hiddenMarkovModel.py:424-486 `add_expectations_file`: 9 transition expectations and the
likelihood line of impl/continuousHmm.c:353-407).
"""
import heapq

import numpy as np


def estimate_band_cells(n_kmers, n_events, expansion):
    """Upper-level estimate of a read's band cells: (lX + lY + 1) diagonals x (expansion + 1) cells, the width
    band_construct gives along dense anchors (reference impl/pairwiseAligner.c:195-246).  Used only to balance shards."""
    return (int(n_kmers) + int(n_events) + 1) * (int(expansion) + 1)


def partition_lpt(work, n_shards):
    """Longest-processing-time-first partition: items sorted by work descending, each dealt to the least-loaded shard.

    Returns a list of n_shards index lists (each in ascending input order, so per-shard outputs keep the caller's
    read order).  Deterministic: ties break on the lower shard id / lower item index.
    """
    work = np.asarray(work, np.int64)
    order = np.lexsort((np.arange(len(work)), -work))
    heap = [(0, s) for s in range(n_shards)]
    shards = [[] for _ in range(n_shards)]
    for i in order:
        load, s = heapq.heappop(heap)
        shards[s].append(int(i))
        heapq.heappush(heap, (load + int(work[i]), s))
    return [sorted(s) for s in shards]


def shard_reads(reads, rank, world, expansion, k):
    """Indices of `reads` (signalalign_b200.Read objects) that `rank` of `world` aligns."""
    work = [estimate_band_cells(max(0, len(r.target) - (k - 1)), r.events.shape[0], expansion) for r in reads]
    return partition_lpt(work, world)[rank]


def expectation_table(result):
    """Per-GPU expectation table of one sa_gpu_expectations_batch result: float64[10] = 9 transition sums
    (from * 3 + to, no pseudocounts) + summed likelihood."""
    t = np.zeros(10)
    if result.transitions is not None and result.n_reads:
        t[:9] = result.transitions.sum(axis=0)
        t[9] = result.likelihood.sum()
    return t


def allreduce_expectations(table, device=None):
    """Sum the per-rank expectation tables over the default process group (the one collective on the path).

    `table`: float64[10 (+ extra slots)] numpy array; returns the reduced numpy array on every rank.  With the NCCL
    backend the buffer lives on `device` (this rank's GPU) and the reduction runs over NVLink / NVSwitch; the message
    is 80 B .. ~1 MB, i.e. latency-bound, so a single ncclAllReduce is the right shape.
    """
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(table, np.float64).copy())
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return t.numpy()
    if dist.get_backend() == "nccl":
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def normalized_transitions(table, n_reads=1, pseudocount=0.001):
    """The M-step the reference trainer applies to the summed table: every `from` row of the 3 x 3 transition
    expectations divided by its sum (reference src/signalalign/hiddenMarkovModel.py:202-211
    normalize_transitions_expectations).  Each per-read expectations file the reference sums starts from `pseudocount`
    in every slot (signalMachine.c:459-470 -> hmm constructor), so n_reads files contribute n_reads * pseudocount."""
    t = np.asarray(table[:9], np.float64).reshape(3, 3) + pseudocount * n_reads
    return t / t.sum(axis=1, keepdims=True)
