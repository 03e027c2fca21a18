"""-m gpu: the collective behind the C-ABI (sa_gpu_comm_*, sa_gpu_allreduce_expectations, sa_gpu_allgather_assignments):
NCCL loaded at run time by libsagpu, no PyTorch involved.  One rank always; two ranks (one process per GPU, the 128-byte
id carried through a file) when the box has two GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

import signalalign_b200 as sa

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import os, sys, time
import numpy as np
sys.path.insert(0, sys.argv[1])
import signalalign_b200 as sa
rank, n, idfile, out = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
sa.gpu_init(rank)
if rank == 0:
    uid = sa.Comm.unique_id()
    with open(idfile + ".tmp", "wb") as f:
        f.write(uid)
    os.rename(idfile + ".tmp", idfile)
else:
    for _ in range(600):
        if os.path.exists(idfile):
            break
        time.sleep(0.1)
    uid = open(idfile, "rb").read()
comm = sa.Comm(uid, n, rank)
table = np.arange(10 + 3 * 4096, dtype=np.float64) * (rank + 1) + 0.25 * rank
red = comm.allreduce_expectations(table.copy())
k = np.arange(5 + 3 * rank, dtype=np.int32) + 100 * rank
e = np.arange(5 + 3 * rank, dtype=np.int32) + 1000 * rank
ak, ae, counts = comm.allgather_assignments(k, e)
np.savez(out, red=red, ak=ak, ae=ae, counts=counts)
comm.close()
"""


def _expected(n):
    base = np.arange(10 + 3 * 4096, dtype=np.float64)
    red = sum(base * (r + 1) + 0.25 * r for r in range(n))
    ak = np.concatenate([np.arange(5 + 3 * r, dtype=np.int32) + 100 * r for r in range(n)])
    ae = np.concatenate([np.arange(5 + 3 * r, dtype=np.int32) + 1000 * r for r in range(n)])
    return red, ak, ae, np.array([5 + 3 * r for r in range(n)], np.int64)


def _run(n, tmp_path):
    idfile = os.path.join(str(tmp_path), "nccl_id_%d" % n)
    procs, outs = [], []
    for r in range(n):
        out = os.path.join(str(tmp_path), "out_%d_%d.npz" % (n, r))
        outs.append(out)
        procs.append(subprocess.Popen([sys.executable, "-c", CHILD, ROOT, str(r), str(n), idfile, out],
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    for p in procs:
        _, err = p.communicate(timeout=600)
        assert p.returncode == 0, err[-2000:]
    red, ak, ae, counts = _expected(n)
    for out in outs:
        with np.load(out) as z:
            assert np.allclose(z["red"], red, rtol=1e-14, atol=0)
            assert np.array_equal(z["ak"], ak) and np.array_equal(z["ae"], ae) and np.array_equal(z["counts"], counts)


def test_comm_single_rank(tmp_path):
    _run(1, tmp_path)


@pytest.mark.skipif(sa.lib().sa_gpu_device_count() < 2, reason="needs two GPUs")
def test_comm_two_ranks_one_process_per_gpu(tmp_path):
    _run(2, tmp_path)
