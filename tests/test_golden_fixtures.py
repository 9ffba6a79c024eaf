"""Frozen vectors of tests/golden/oracle_r01.npz (made by tests/golden/make_fixtures.py from the CPU
oracle; seeded inputs are rebuilt by tests/golden/cases.py):
  * CPU: the oracle still reproduces them (guards the checker itself);
  * GPU: the CUDA path reproduces them through the C ABI, bit for bit.
The reference's own known answers (ref_dist_kat.json) are tested in test_oracle_golden.py (oracle)
and test_gpu_parity.py::test_sad_satd_reference_goldens (CUDA)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import cases  # noqa: E402


@pytest.fixture(scope="module")
def frozen():
    return dict(np.load(os.path.join(HERE, "golden", "oracle_r01.npz")))


def _compare(got, frozen):
    assert set(got) == set(frozen)
    bad = [k for k in sorted(frozen)
           if not np.array_equal(np.asarray(got[k]).astype(frozen[k].dtype).reshape(frozen[k].shape), frozen[k])]
    assert not bad, bad


def test_oracle_reproduces_frozen_vectors(oracle, frozen):
    _compare(cases.oracle_vectors(oracle), frozen)


def test_fixture_covers_every_kernel_class(frozen):
    for bd in (8, 10, 12):
        for k in ("sad", "satd", "mc_put", "mc_prep", "mc_avg", "fht", "sse", "cdef"):
            assert "%s_%d" % (k, bd) in frozen
        assert frozen["fht_%d" % bd].size == 193  # every valid (TxSize, TxType)
    assert frozen["frame_cost_8"].size == frozen["frame_mv_8"].shape[0] > 1000


@pytest.mark.gpu
def test_cuda_reproduces_frozen_vectors(oracle, frozen):
    import torch
    import rav1e_b200 as r
    assert torch.cuda.is_available()
    ctx = r.Context(0)
    try:
        _compare(cases.cuda_vectors(r, ctx, oracle), frozen)
    finally:
        ctx.close()
