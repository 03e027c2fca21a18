#!/usr/bin/env python
"""bench.py — events aligned/s and banded-DP GCUPS of the signalAlign hot path on B200 (BASELINE.json metric).

A "step" = one pass of the hot path (a batch of getAlignedPairsUsingAnchors calls) over one batch of synthetic reads
of BASELINE config 3's shape: ~10k events per read against a synthetic reference, 6-mer Gaussian table, band
expansion 100, threshold 0.01 (signalalign_b200/synth.py, seed fixed).  Per-GPU work is fixed (weak scaling): every
rank aligns its own shard of reads, no data-path collective.

    value   events/s with the batch already resident in HBM (sa_batch_run only), device-timed with CUDA events
    e2e     the same through the host-buffer C-ABI call (packing + H2D + kernels + D2H inside the timed region)
    roofline / roofline_compute / cpu_baseline: see DESIGN.md §Measurement

`--impl reference` times the reference's own CPU implementation (oracle/_ref, built from /root/reference by
oracle/Makefile; else the oracle port) on the box's host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = ("config3-synthetic: ~10k-event reads vs synthetic reference, ACGT 6-mer Gaussian table, -x 100, threshold 0.01, "
            "guide anchors = match runs of geometric length 1/0.12 trimmed by 14 (SURVEY 8d)")
EXPANSION, TRACE_BACK, MIN_DIAGS, THRESHOLD = 100, 50, 1000, 0.01
# guide anchors: SURVEY.md 8(d)'s generator (signalalign_b200/synth.py ANCHOR_PATTERNS); --anchors picks another pattern
ANCHOR_PATTERN = "contract"
CPU_EXPECTATIONS = False     # --workload config4: the CPU arm times getExpectationsUsingAnchors
from signalalign_b200.synth import ANCHOR_PATTERNS  # noqa: E402  (numpy only)
GEN = dict(ANCHOR_PATTERNS[ANCHOR_PATTERN])


def compare_pairs(ra, rb):
    """posterior agreement of two results on the same reads (the two precisions): pairs matched on (read, x, y)"""
    def keys(r):
        n = np.diff(r.pair_offsets)
        rid = np.repeat(np.arange(len(n), dtype=np.int64), n)
        return (rid << 42) | (r.pairs[:, 1].astype(np.int64) << 21) | r.pairs[:, 2].astype(np.int64)
    ka, kb = keys(ra), keys(rb)
    if ka.shape == kb.shape and np.array_equal(ka, kb):
        ia = ib = slice(None)
        only = np.zeros(0, dtype=np.int64)
    else:
        _, ia, ib = np.intersect1d(ka, kb, assume_unique=True, return_indices=True)
        only = np.concatenate([np.delete(ra.pairs[:, 0], ia), np.delete(rb.pairs[:, 0], ib)]).astype(np.int64)
    dp = np.abs(ra.pairs[ia, 0].astype(np.int64) - rb.pairs[ib, 0].astype(np.int64))
    return {"pairs_compared": int(dp.shape[0]), "max_abs_dp": float(dp.max()) / 1e7 if dp.size else 0.0,
            "dp_above_1e-4": int((dp > 1000).sum()), "dp_above_1e-5": int((dp > 100).sum()),
            "kmer_mismatches": int((ra.pairs[ia, 3] != rb.pairs[ib, 3]).sum()),
            "pairs_in_one_only": int(only.shape[0]), "largest_p_of_those": float(only.max()) / 1e7 if only.size else None}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=int(os.environ.get("SA_BENCH_READS", "4736")),
                    help="reads per GPU per step (default 32 per SM: ~6 work items per resident CTA, so the tail of the "
                         "persistent grid is a few % of the step)")
    ap.add_argument("--cpu-reads", type=int, default=0, help="reads in the CPU sample (default: one per host core)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-expectations", action="store_true", help="skip the config-4 (expectations) figure")
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--anchors", default=ANCHOR_PATTERN, choices=sorted(ANCHOR_PATTERNS),
                    help="guide-anchor pattern of the synthetic reads (default: SURVEY 8d's)")
    ap.add_argument("--precision", default="fp64", choices=["fp32", "fp64"],
                    help="arithmetic of the band kernel on the alignment path (config 3): fp64 = the reference's double arithmetic "
                         "bit for bit (the headline: every posterior within the contract's 1e-4); fp32 = DP state and logAdd in "
                         "single precision relative to per-diagonal integer offsets (sa_gpu_set_precision): 1.45x faster, but ~1 "
                         "posterior in a million leaves the 1e-4 band where a float decides the reference's 7.5 cut-off the other "
                         "way (measured in every run: config.other_precision_rank0).  Expectations and HDP runs are always fp64")
    ap.add_argument("--no-other-precision", action="store_true", help="skip the secondary figure in the other precision")
    ap.add_argument("--no-dense", action="store_true", help="skip the secondary figure on round 1's dense anchor pattern")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: ONE read set of --reads reads (same seed on every rank) split across the ranks by "
                         "signalalign_b200.shard.partition_lpt (longest-processing-time-first on estimated band cells); the "
                         "default is weak scaling (every rank its own --reads reads)")
    ap.add_argument("--workload", default="config3", choices=["config2", "config3", "config4", "config5"],
                    help="config3 (default, BASELINE's headline): alignment of ~10k-event reads, 6-mer Gaussian table; config4: "
                         "the expectations pass (getExpectationsUsingAnchors) on the same reads; config5: 5-mer table, HDP "
                         "emission (synthetic spline tables), log-normal read lengths (median ~20k events, tail to 400k)")
    a = ap.parse_args()
    global CPU_EXPECTATIONS
    CPU_EXPECTATIONS = a.workload == "config4"
    GEN.clear()
    GEN.update(ANCHOR_PATTERNS[a.anchors])
    return a


# ------------------------------------------------------------------------------------------------ CPU arms
def write_model_file(path, table, alphabet="ACGT", k=6):
    """Synthetic table in the reference's .model text format (stateMachine.c:1445-1452)."""
    t = [0.79, 0.1965, 0.0133, 0.8035, 0.1965, 0.0, 0.9867, 1e-9, 0.0133]
    with open(path, "w") as f:
        f.write(f"3\t{len(alphabet)}\t{alphabet}\t{k}\n")
        f.write("\t".join(repr(v) for v in t) + "\t0.0\n")
        f.write("\t".join(repr(float(v)) for v in table.reshape(-1)) + "\n")


def _cpu_worker(args):
    kind, model_path, reads = args
    cells = 0
    t0 = time.time()
    if kind == "reference":
        from oracle import ref
        sm = ref.RefStateMachine(model_path, kind=0)
        for r in reads:
            sm.set_scaling(r["scale"], r["shift"], r["var"])
            fn = sm.expectations if CPU_EXPECTATIONS else sm.align
            fn(r["target"], r["events"], r["anchors"], threshold=THRESHOLD, min_diags=MIN_DIAGS,
               trace_back=TRACE_BACK, expansion=EXPANSION)
    else:
        from oracle import port
        pm = port.load_model_file(model_path)
        p = port.params(THRESHOLD, MIN_DIAGS, TRACE_BACK, EXPANSION)
        for r in reads:
            pm.scale, pm.shift, pm.var = r["scale"], r["shift"], r["var"]
            _, st = port.align(pm, r["target"], r["events"], r["anchors"], p)
            cells += st[0]
    return time.time() - t0, cells


def cpu_arm(reads, table, cores):
    """Time the reference CPU path (process per core, like signalAlign's own per-read process model)."""
    import multiprocessing as mp
    from oracle import ref
    kind = "reference" if ref.available() else "port"
    tmp = tempfile.mkdtemp(prefix="sa_bench_")
    model_path = os.path.join(tmp, "synthetic.model")
    write_model_file(model_path, table)
    shards = [reads[i::cores] for i in range(cores)]
    shards = [s for s in shards if s]
    ctx = mp.get_context("fork")
    t0 = time.time()
    with ctx.Pool(len(shards)) as pool:
        pool.map(_cpu_worker, [(kind, model_path, s) for s in shards])
    wall = time.time() - t0
    events = sum(r["events"].shape[0] for r in reads)
    return kind, wall, events, len(shards)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ main
def main():
    a = parse()
    # stdout carries the ONE JSON line and nothing else: library banners (NCCL prints its version there) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(obj), flush=True)
        os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from signalalign_b200 import synth

    if a.impl == "reference":
        if rank != 0:
            return
        cores = os.cpu_count() or 1
        n = a.cpu_reads or cores
        reads, table = synth.make_batch(n, seed=a.seed, **GEN)
        walls = []
        kind = "reference"
        for i in range(a.warmup + a.steps):
            k = 1 if i < a.warmup else None
            sample = reads[:max(1, min(len(reads), cores))] if k else reads
            kind, wall, events, used = cpu_arm(sample, table, cores)
            if i >= a.warmup:
                walls.append((wall, events, used))
        wall = sum(w for w, _, _ in walls)
        events = sum(e for _, e, _ in walls)
        used = walls[0][2]
        v = events / wall
        emit(({
            "impl": "reference", "metric": "events_aligned_per_sec", "value": v, "unit": "events/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * wall / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "reads_per_step": len(reads)},
            "cpu_baseline": {"value": v, "unit": "events/s", "cores": used, "kind": kind,
                             "sample": f"{len(reads)} reads per step, one process per core"},
            "e2e": {"value": v, "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    import signalalign_b200 as sa

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    sa.gpu_init(local_rank)
    # every rank gets its share of the host cores for packing (torchrun exports OMP_NUM_THREADS=1)
    host_threads = max(1, (os.cpu_count() or 1) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
    sa.set_host_threads(host_threads)

    trans = synth.nanopore_transitions()
    expect = a.workload == "config4"
    if a.workload == "config5":
        # RNA-like: ~1.3 events per nt, median 20 k events; the RNA .nhdp of BASELINE config 5 is absent from the reference
        # snapshot: spline tables of the same shape (one row per kmer, 1200 grid points, two-component mixtures)
        if a.reads == 4736:
            a.reads = 3072
        reads_d, table = synth.make_batch(a.reads, seed=a.seed + 1 + 1000 * rank, alphabet="ACGT", k=5, ref_len=3_000_000,
                                          mean_len=15400, sigma_len=1.35, min_len=770, max_len=300000,
                                          n_probs=(0.05, 0.70, 0.18, 0.07), **GEN)
        model = sa.Model.from_tables("ACGT", 5, trans, table, kind=sa.EMIT_HDP)
        grid = np.linspace(50.0, 140.0, 1200)
        mu = table[:, 0]
        pp = 0.7 * np.exp(-0.5 * ((grid[None, :] - mu[:, None]) / 2.0) ** 2) / (2.0 * np.sqrt(2 * np.pi)) + \
            0.3 * np.exp(-0.5 * ((grid[None, :] - mu[:, None] - 3.0) / 4.0) ** 2) / (4.0 * np.sqrt(2 * np.pi))
        model.set_hdp(grid, np.arange(table.shape[0], dtype=np.int32), pp, np.gradient(pp, grid, axis=1))
        a.no_cpu = a.no_dense = a.no_expectations = True
    elif a.workload == "config2":
        # cytosine-variant calling: ACEGT 6-mer table (E = 5mC: the C kmer's level shifted by 1.5 pA per E), config 3's reads,
        # 10 % of the reference's C replaced by the two-way ambiguity code P (C/E, create_ambig_bases
        # impl/pairwiseAligner.c:32-65) -> multi-path cells (general kernel)
        if a.reads == 4736:
            a.reads = 592
        reads_d, t4 = synth.make_batch(a.reads, seed=a.seed + 2 + 1000 * rank, **GEN)
        K5 = 5 ** 6
        digits = np.zeros((K5, 6), np.int64)
        v = np.arange(K5)
        for j in range(5, -1, -1):
            digits[:, j] = v % 5
            v //= 5
        to4 = np.array([0, 1, 1, 2, 3])
        twin = np.zeros(K5, np.int64)
        for j in range(6):
            twin = twin * 4 + to4[digits[:, j]]
        table = t4[twin].copy()
        table[:, 0] += 1.5 * (digits == 2).sum(1)
        model = sa.Model.from_tables("ACEGT", 6, trans, table)
        rng2 = np.random.default_rng(a.seed + 11 + rank)
        for r in reads_d:
            t = np.frombuffer(r["target"].encode(), np.uint8).copy()
            c = np.nonzero(t == ord("C"))[0]
            t[c[rng2.random(len(c)) < 0.1]] = ord("P")
            r["target"] = t.tobytes().decode()
        a.no_cpu = a.no_dense = a.no_expectations = True
    else:
        reads_d, table = synth.make_batch(a.reads, seed=a.seed + (0 if a.strong else 1000 * rank), **GEN)
        model = sa.Model.from_tables("ACGT", 6, trans, table)
    shard_info = None
    if a.strong and world > 1:
        from signalalign_b200 import shard
        work = [shard.estimate_band_cells(max(0, len(r["target"]) - 5), r["events"].shape[0], EXPANSION) for r in reads_d]
        parts = shard.partition_lpt(work, world)
        loads = [int(sum(work[i] for i in part)) for part in parts]
        shard_info = {"reads_total": len(reads_d), "reads_per_rank": [len(q) for q in parts],
                      "estimated_load_imbalance": max(loads) / (sum(loads) / world)}
        reads_d = [reads_d[i] for i in parts[rank]]
    if expect:
        a.no_dense = a.no_expectations = True
    prec = a.precision if a.workload == "config3" else "fp64"
    sa.set_precision(prec)
    p = sa.default_params(threshold=THRESHOLD, expansion=EXPANSION, trace_back=TRACE_BACK, min_diags=MIN_DIAGS)
    reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
    events = sum(r.events.shape[0] for r in reads)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # L2 flush buffer (126 MB L2): written between timed steps; the batch itself is also far larger than L2
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    # ---- resident arm: inputs in HBM before the timed region
    batch = sa.Batch(model, reads, p, expectations=expect)
    cells = batch.band_cells
    batch_h2d = batch.h2d_bytes
    kernel_items, wide_items = batch.kernel_items, batch.wide_items
    n_diag = sum(len(r.target) - 5 + r.events.shape[0] + 1 for r in reads)
    for _ in range(a.warmup):
        batch.run()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev_ms, kern_ms, all_ms, launches = [], [], [], 0
    barrier()
    t0 = time.time()
    for _ in range(a.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        batch.run()
        ms, n = batch.timing()
        dev_ms.append(sum(ms))
        kern_ms.append(ms[0])
        all_ms.append(ms)
        launches += n
    barrier()
    wall_resident = time.time() - t0
    clocks = sampler.stop() if rank == 0 else None
    step_ms = float(np.sum(dev_ms)) / a.steps
    res = batch.fetch()
    res_keep = res if (a.workload == "config3" and not a.no_other_precision) else None
    n_pairs = int(res.pair_offsets[-1])
    if expect:        # the expectations pass returns 9 transition sums + the likelihood per read
        pairs_digest = [float(v) for v in res.transitions.sum(0)] + [float(res.likelihood.sum())]
    else:
        pairs_digest = [int(res.pairs[:, c].astype(np.int64).sum()) for c in range(4)]   # p, x, y, kmer column sums
    batch.close()

    # ---- secondary, labelled: the same batch, resident, in the other precision
    other = None
    if a.workload == "config3" and not a.no_other_precision:
        oprec = "fp64" if prec == "fp32" else "fp32"
        sa.set_precision(oprec)
        ob = sa.Batch(model, reads, p)
        ob.run()
        oms = []
        for _ in range(2):
            flush.fill_(1)
            torch.cuda.synchronize()
            ob.run()
            oms.append(sum(ob.timing()[0]))
        ores = ob.fetch()
        other = {"precision": oprec, "gcups": ob.band_cells / (min(oms) * 1e-3) / 1e9, "events_per_s": events / (min(oms) * 1e-3),
                 "ms_per_step": min(oms), "aligned_pairs": int(ores.pair_offsets[-1]),
                 "pairs_column_sums": [int(ores.pairs[:, c].astype(np.int64).sum()) for c in range(4)],
                 "agreement_with_headline": compare_pairs(res_keep, ores)}
        del ores
        res_keep = None
        ob.close()
        sa.set_precision(prec)

    # ---- secondary, labelled: the same resident measurement on round 1's dense anchor pattern (every kmer an anchor)
    dense = None
    if not a.no_dense and a.anchors != "dense":
        dn = max(148, a.reads // 4)
        dreads_d, _ = synth.make_batch(dn, seed=a.seed + 1000 * rank + 7, table=table, **ANCHOR_PATTERNS["dense"])
        dreads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in dreads_d]
        db = sa.Batch(model, dreads, p)
        db.run()
        dms = []
        for _ in range(2):
            flush.fill_(1)
            torch.cuda.synchronize()
            db.run()
            dms.append(sum(db.timing()[0]))
        dense = {"anchor_pattern": "dense (trim 0, runs of ~80): round 1's workload", "reads": dn,
                 "gcups": db.band_cells / (min(dms) * 1e-3) / 1e9,
                 "events_per_s": sum(r.events.shape[0] for r in dreads) / (min(dms) * 1e-3),
                 "work_items_by_kernel": dict(zip(("general", "band_32_64", "band_128_256", "wide"), db.kernel_items))}
        db.close()
        del dreads, dreads_d

    # ---- end-to-end arm: host buffers -> C-ABI call -> host result (pack + H2D + kernels + D2H timed)
    e2e_t = []
    h2d = d2h = 0
    barrier()
    for i in range(max(2, a.steps) + 1):           # one untimed call (pinned / device pools), then the timed ones
        t1 = time.time()
        # sa_gpu_align_batch / sa_gpu_expectations_batch: host buffers in, host result out
        r2 = sa.expectations_batch(model, reads, p) if expect else sa.align_batch(model, reads, p)
        if i > 0:
            e2e_t.append(time.time() - t1)
        assert int(r2.pair_offsets[-1]) == n_pairs, "the one-shot call and the staged batch disagree"
        if expect:
            assert np.allclose(r2.transitions.sum(0), pairs_digest[:9], rtol=1e-9)
            h2d, d2h = batch_h2d, 10 * 8 * len(reads)
        else:
            h2d, d2h, _ = sa.last_call_stats()
        del r2
    barrier()
    e2e_s = float(np.mean(e2e_t))

    # ---- config 4 beside it: getExpectationsUsingAnchors on the same reads, end to end (one warm-up, one timed call)
    exp_s = None
    if not a.no_expectations:
        from signalalign_b200 import shard
        exp_table = exp_ar_ms = None
        for i in range(2):
            t1 = time.time()
            r3 = sa.expectations_batch(model, reads, p)
            exp_s = time.time() - t1
            exp_table = shard.expectation_table(r3)
            del r3
        barrier()
        # the one collective of the path: all-reduce of the per-GPU expectation tables through the C-ABI
        # (sa_gpu_allreduce_expectations: libsagpu's own NCCL communicator over NVLink at N > 1; torch.distributed only
        # carries the 128-byte id to the ranks), checked against the all-gathered per-rank tables
        if world > 1:
            ids = [sa.Comm.unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            comm = sa.Comm(ids[0], world, rank)
            comm.allreduce_expectations(exp_table.copy())                                        # warm-up (communicator)
            barrier()
            t1 = time.time()
            reduced = comm.allreduce_expectations(exp_table.copy())
            exp_ar_ms = 1e3 * (time.time() - t1)
            parts = [torch.zeros(10, dtype=torch.float64, device="cuda") for _ in range(world)]
            dist.all_gather(parts, torch.from_numpy(exp_table).to("cuda"))
            want = torch.stack(parts).sum(0).cpu().numpy()
            assert np.allclose(reduced, want, rtol=1e-12, atol=0), (reduced, want)
            exp_table = reduced
            comm.close()

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    step_ms_max = reduce_max(step_ms)
    kern_ms_max = reduce_max(float(np.mean(kern_ms)))
    e2e_max = reduce_max(e2e_s)
    events_all = reduce_sum(float(events))
    cells_all = reduce_sum(float(cells))

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        # algorithmic bytes per band cell of the DP kernel: forward match plane written + read (16 B), X/Y
        # planes on 2 of 10 rows (6.4 B), inputs / outputs 0.4 B
        alg_bytes_per_cell = 22.8 if prec == "fp64" else 11.6      # (fp32: 4-byte planes)
        achieved_gbs = cells * alg_bytes_per_cell / (float(np.mean(kern_ms)) * 1e-3) / 1e9
        # measured DRAM traffic of the kernel (one ncu --set full capture, profiles/): bytes per band cell x this launch's cells
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r2_dram_per_cell.json")))[a.anchors][prec]["bytes_per_cell"] * cells
        except Exception:
            pass
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        # FP64 pipe: ~150 algorithmic FP64 lane-instructions per band cell; peak = measured DFMA/DADD/DMUL
        # issue rate of this pool's B200 (tools/peak_fp.cu -> profiles/r1_peak_fp.jsonl), else 64 lanes/SM/clk nominal
        fp64_per_cell = 150.0
        # FP32 mode: 122 algorithmic FP32 lane-instructions per band cell (forward 53, backward 67,
        # posterior 2); peak = measured FFMA issue rate
        fp32_per_cell = 122.0
        fp32_peak, fp32_src = 148 * 128 * 1.965e9, "nominal 128 lanes/SM/clk at 1965 MHz"
        fp64_peak, fp64_src = 148 * 64 * 1.965e9, "nominal 64 lanes/SM/clk at 1965 MHz"
        try:
            for line in open(os.path.join(ROOT, "profiles", "r1_peak_fp.jsonl")):
                rec = json.loads(line)
                if rec.get("pipe") == "dadd":
                    fp64_peak, fp64_src = rec["lane_instr_per_s"], "measured (profiles/r1_peak_fp.jsonl, tools/peak_fp.cu)"
                if rec.get("pipe") == "ffma":
                    fp32_peak, fp32_src = rec["lane_instr_per_s"], "measured (profiles/r1_peak_fp.jsonl, tools/peak_fp.cu)"
        except Exception:
            pass
        out = {
            "metric": "events_aligned_per_sec", "value": events_all / (step_ms_max * 1e-3), "unit": "events/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": step_ms_max,
            "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None,
            "dtype": "f32" if prec == "fp32" else "f64", "data": "synthetic",
            "config": {"strong_scaling_shards": shard_info,
                       "workload": {"config2": "config2-synthetic: ACEGT 6-mer table, config 3's reads with 10 % of the reference's C "
                                               "flagged as the C/E ambiguity code P (multi-path cells, general kernel)",
                                    "config3": WORKLOAD if a.anchors == ANCHOR_PATTERN else "config3-synthetic, anchor pattern " + a.anchors,
                                    "config4": "config4-synthetic: expectations pass (getExpectationsUsingAnchors) over config 3's reads",
                                    "config5": "config5-synthetic: ACGT 5-mer table, HDP emission (1024 rows x 1200 grid points), "
                                               "log-normal read lengths (median ~20k events, clipped at 400k), -x 100"}[a.workload],
                       "anchor_pattern": {"name": a.anchors, **ANCHOR_PATTERNS[a.anchors]},
                       "precision": {"fp32": "fp32: DP state, stored rows and logAdd in single precision relative to per-diagonal "
                                             "integer offsets; emission differences, offsets and total probability in double "
                                             "(opt-in fast mode: see other_precision_rank0.agreement_with_headline)",
                                     "fp64": "fp64: the reference's double arithmetic, bit for bit"}[prec],
                       "other_precision_rank0": other,
                       "mean_band_width_rank0": cells / n_diag, "dense_pattern_rank0": dense,
                       "reads_per_gpu_per_step": a.reads, "host_threads_per_rank": host_threads, "events_per_step": events_all,
                       "band_cells_per_step": cells_all, "l2": "flushed between steps (256 MiB write) and inputs > L2",
                       "aligned_pairs_rank0": n_pairs,
                       ("expectation_sums_rank0" if expect else "pairs_column_sums_rank0"): pairs_digest,
                       "events_per_read_rank0": {"median": float(np.median([r.events.shape[0] for r in reads])),
                                                 "max": int(max(r.events.shape[0] for r in reads))},
                       "expectations_e2e_rank0": None if exp_s is None else {
                           "call": "sa_gpu_expectations_batch (config 4 shape: same reads)", "events_per_s": events / exp_s,
                           "gcups": cells / exp_s / 1e9, "ms": 1e3 * exp_s,
                           "allreduce_ms": exp_ar_ms, "table_all_ranks": [float(v) for v in exp_table]},
                       "kernel_ms_rank0": {"wide_and_band_128_256_slots": float(np.mean([m[0] for m in all_ms])),
                                           "band_32_64_slots": float(np.mean([m[1] for m in all_ms])),
                                           "general": float(np.mean([m[2] for m in all_ms]))},
                       "work_items_by_kernel_rank0": {"general": kernel_items[0], "band_32_64_slots": kernel_items[1],
                                                      "band_128_256_slots": kernel_items[2], "wide": kernel_items[3],
                                                      "wide_by_window": dict(zip(("256", "512", "1024", "2048", "4096"), wide_items))}},
            "gcups": cells_all / (step_ms_max * 1e-3) / 1e9,
            "gpu_launches": launches,
            "clocks": clocks,
            "e2e": {"value": events_all / e2e_max, "unit": "events/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "gcups": cells_all / e2e_max / 1e9,
                    "ms_each_call_rank0": [round(1e3 * t, 1) for t in e2e_t]},
            # the binding roofline of the band kernels is the FP64 pipe; HBM is the secondary object
            "roofline": ({"bound": "fp32_pipe", "achieved": cells * fp32_per_cell / (float(np.mean(kern_ms)) * 1e-3),
                          "peak": fp32_peak, "peak_source": fp32_src, "unit": "FP32 lane-instr/s",
                          "frac": cells * fp32_per_cell / (float(np.mean(kern_ms)) * 1e-3) / fp32_peak,
                          "per_cell": fp32_per_cell, "clock_mhz": sm_mhz, "traffic": traffic,
                          "note": "not a contraction (no tensor cores) and not HBM bound: single-precision cubic logAdd; the "
                                  "kernel is issue bound (FP32 is a quarter of its instruction mix, profiles/), see roofline_hbm"}
                         if prec == "fp32" else
                         {"bound": "fp64_pipe", "achieved": cells * fp64_per_cell / (float(np.mean(kern_ms)) * 1e-3),
                          "peak": fp64_peak, "peak_source": fp64_src, "unit": "FP64 lane-instr/s",
                          "frac": cells * fp64_per_cell / (float(np.mean(kern_ms)) * 1e-3) / fp64_peak,
                          "per_cell": fp64_per_cell, "clock_mhz": sm_mhz, "traffic": traffic,
                          "note": "not a contraction (no tensor cores) and not HBM bound: FP64 cubic logAdd, see roofline_hbm"}),
            "roofline_hbm": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                             "frac": achieved_gbs / hbm_peak, "traffic": traffic, "algorithmic_bytes": alg_bytes_per_cell * cells,
                             "peak_source": peak_src},
            "wall_s_resident_loop": wall_resident,
        }
        if world == 1 and not a.no_cpu:
            cores = os.cpu_count() or 1
            n = a.cpu_reads or min(cores, len(reads_d))
            kind, wall, ev, used = cpu_arm(reads_d[:n], table, cores)
            out["cpu_baseline"] = {"value": ev / wall, "unit": "events/s", "cores": used, "kind": kind,
                                   "sample": f"first {n} reads of the step's batch, one process per core, {wall:.1f} s"}
        emit(out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
