"""-m gpu: the CUDA path, called through the C-ABI, against the oracle on the same seeded inputs."""
import numpy as np
import pytest

import signalalign_b200 as sa
from oracle import port
from signalalign_b200 import synth

from helpers import compare_pairs, oracle_align, port_model_from, port_params, sa_read

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small_batch():
    reads, table = synth.make_batch(6, seed=11, mean_len=500, min_len=200, max_len=900)
    return reads, table


def test_gaussian_meanonly_small(small_batch):
    reads, table = small_batch
    trans = synth.nanopore_transitions()
    model = sa.Model.from_tables("ACGT", 6, trans, table)
    pm = port_model_from("ACGT", 6, trans, table)
    p = sa.default_params(threshold=0.01, expansion=50, trace_back=50, min_diags=200)
    res = sa.align_batch(model, [sa_read(r) for r in reads], p)
    worst = 0
    for i, r in enumerate(reads):
        ora, stats = oracle_align(pm, r, port_params(p))
        worst = max(worst, compare_pairs(res.read_pairs(i), ora, p.threshold))
    assert worst <= 10  # FP64 device path: differences only from exp/log last-bit effects
