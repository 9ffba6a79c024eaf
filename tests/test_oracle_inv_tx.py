"""CPU: the oracle's inverse transforms (oracle/r1o_itx.c) against
  * tests/golden/inv_tx_r01.npz: outputs of the reference's own functions (tests/golden/make_inv_tx_golden.py);
  * the reference's `roundtrips` test (src/transform/mod.rs:589-720): forward + inverse gives the source
    back within the listed tolerance, with its 37 (size, type, tolerance) rows."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(HERE, "golden", "inv_tx_r01.npz")))


def test_golden_has_every_valid_pair(oracle, golden):
    pairs = {(int(k.split("_")[1]), int(k.split("_")[2])) for k in golden if k.startswith("c_")}
    want = {(s, t) for s in range(19) for t in range(16) if oracle.inv_txfm_valid(s, t)}
    assert pairs == want and len(want) == 193


def test_inverse_2d_reproduces_reference_vectors(oracle, golden):
    for k in sorted(golden):
        if not k.startswith("c_"):
            continue
        _, s, t, bd = k.split("_")
        got = oracle.inv_txfm2d(golden[k], int(s), int(t), int(bd))
        assert np.array_equal(got, golden["r" + k[1:]]), k


def test_inverse_add_reproduces_reference_vectors(oracle, golden):
    n = 0
    for k in sorted(golden):
        if not k.startswith("ac_"):
            continue
        _, s, t, bd = k.split("_")
        plane = golden["d" + k[2:]].copy()
        oracle.inv_txfm_add_batch(golden[k], plane, [[0, 0]], int(s), int(t), int(bd))
        assert np.array_equal(plane, golden["o" + k[2:]]), k
        n += 1
    assert n == 8


def test_forward_2d_reproduces_reference_vectors(oracle):
    """tests/golden/fwd_tx_r01.npz: residual -> coefficients made by the reference's own forward 1-D functions
    (tests/golden/make_fwd_tx_golden.py), all 193 pairs"""
    g = dict(np.load(os.path.join(HERE, "golden", "fwd_tx_r01.npz")))
    n = 0
    for k in sorted(g):
        if not k.startswith("x_"):
            continue
        s, t, bd = (int(v) for v in k.split("_")[1:])
        assert np.array_equal(oracle.fht(g[k], s, t, bd), g["y" + k[1:]]), k
        n += 1
    assert n == 193


# (tx_size, tx_type, tolerance): src/transform/mod.rs:665-708; sizes / types by their enum values
ROUNDTRIPS = ([(0, t, 0) for t in (0, 1, 2, 3, 9, 10, 11, 12, 13)]
              + [(1, 0, 1), (1, 1, 1), (1, 2, 1), (1, 3, 1), (1, 9, 0), (1, 10, 0), (1, 11, 0), (1, 12, 0), (1, 13, 1)]
              + [(2, 0, 1), (2, 1, 1), (2, 2, 1), (2, 3, 1), (2, 9, 0), (2, 10, 1), (2, 11, 1)]
              + [(3, 0, 2), (3, 9, 0)]
              + [(5, 0, 1), (6, 0, 1), (13, 0, 1), (14, 0, 1), (7, 0, 1), (8, 0, 1), (15, 0, 2), (16, 0, 2),
                 (9, 0, 2), (10, 0, 2)])


@pytest.mark.parametrize("tx_size,tx_type,tol", ROUNDTRIPS)
def test_reference_roundtrips(oracle, tx_size, tx_type, tol):
    rng = np.random.default_rng(100 * tx_size + tx_type)
    w, h = oracle.lib().r1o_tx_width(tx_size), oracle.lib().r1o_tx_height(tx_size)
    for _ in range(8):
        src = rng.integers(0, 256, (h, w))
        dst = rng.integers(0, 256, (h, w)).astype(np.uint8)
        freq = oracle.fht((src - dst).astype(np.int16), tx_size, tx_type, 8)
        oracle.inv_txfm_add_batch(freq.reshape(-1), dst, [[0, 0]], tx_size, tx_type, 8)
        assert np.abs(src - dst.astype(np.int64)).max() <= tol
