"""CPU suite, part 4: the N>1 plumbing with world_size 2 over gloo — read sharding and the one collective of the path
(all-reduce of the per-GPU expectation tables).  No GPU involved: the per-rank tables are synthetic."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

import signalalign_b200 as sa
from signalalign_b200 import shard, synth


def test_partition_lpt_balances_and_covers():
    rng = np.random.default_rng(2)
    work = np.round(rng.lognormal(np.log(6000), 0.8, 500)).astype(np.int64)
    for n in (1, 2, 4, 8):
        parts = shard.partition_lpt(work, n)
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(500))                       # every read exactly once
        loads = np.array([work[p].sum() for p in parts])
        assert loads.max() - loads.min() <= work.max()        # LPT bound
        assert all(p == sorted(p) for p in parts)             # input order preserved inside a shard
    assert shard.partition_lpt([], 4) == [[], [], [], []]
    assert shard.partition_lpt([5], 2) == [[0], []]


def test_shard_reads_uses_band_cells():
    reads_d, _ = synth.make_batch(12, seed=4, mean_len=300, min_len=100, max_len=900)
    reads = [sa.Read(r["target"], r["events"], r["anchors"]) for r in reads_d]
    a, b = shard.shard_reads(reads, 0, 2, 100, 6), shard.shard_reads(reads, 1, 2, 100, 6)
    assert sorted(a + b) == list(range(12)) and not set(a) & set(b)


def _worker(rank, world, port_no, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    table = np.concatenate([rng.uniform(0, 50, 9), [-rng.uniform(1e5, 1e6)]])
    out = shard.allreduce_expectations(table)
    q.put((rank, table, out))
    dist.barrier()
    dist.destroy_process_group()


def test_allreduce_expectations_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port_no = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port_no, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    total = res[0][1] + res[1][1]
    for _, _, out in res:
        assert np.allclose(out, total, rtol=1e-15, atol=0)
    t = shard.normalized_transitions(total, n_reads=2)
    assert np.allclose(t.sum(axis=1), 1.0)


def test_allreduce_is_identity_without_process_group():
    t = np.arange(10, dtype=np.float64)
    assert np.array_equal(shard.allreduce_expectations(t), t)
