"""GPU parity of the inverse transforms (csrc/itx*.cu*) through the C ABI:
  * against tests/golden/inv_tx_r01.npz -- outputs of the reference's own functions
    (tests/golden/make_inv_tx_golden.py), all 193 (TxSize, TxType) pairs;
  * against the oracle (oracle/r1o_itx.c) on whole planes of random non-overlapping blocks, 3 bit depths;
  * the reference's forward + inverse roundtrip (src/transform/mod.rs:589-720) run entirely on the device."""
import os

import numpy as np
import pytest

from gpu_util import twin_plane

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
BD = [(1, 8), (2, 10), (2, 12)]


@pytest.fixture(scope="module")
def r():
    import rav1e_b200 as mod
    return mod


@pytest.fixture(scope="module")
def ctx(r):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    c = r.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(HERE, "golden", "inv_tx_r01.npz")))


def _grid_jobs(r, rng, n, w, h, W, H, x0=0, y0=0):
    """n distinct cells of the w x h grid over a W x H area, in random order"""
    cells = [(x, y) for y in range(y0, y0 + H - h + 1, h) for x in range(x0, x0 + W - w + 1, w)]
    pick = rng.permutation(len(cells))[:n]
    jobs = np.zeros(len(pick), dtype=r.ITX_JOB)
    for i, k in enumerate(pick):
        jobs[i] = cells[k]
    return jobs


def test_inverse_reference_vectors(r, ctx, oracle, golden):
    """the reference's residual for every valid pair, added to a random plane"""
    rng = np.random.default_rng(5)
    planes = {bd: twin_plane(r, oracle, ctx, 96, 80, pb, bd, rng, xpad=16, ypad=16) for pb, bd in BD}
    n = 0
    for k in sorted(golden):
        if not k.startswith("c_"):
            continue
        s, t, bd = (int(v) for v in k.split("_")[1:])
        o, d = planes[bd]
        res = golden["r" + k[1:]]
        h, w = res.shape
        x, y = int(rng.integers(0, 96 - w + 1)), int(rng.integers(0, 80 - h + 1))
        want = o.visible[y:y + h, x:x + w].astype(np.int64) + res
        o.visible[y:y + h, x:x + w] = np.clip(want, 0, (1 << bd) - 1).astype(o.dtype)
        r.inverse_transform_add(ctx, golden[k], d, np.array([(x, y)], dtype=r.ITX_JOB), s, t, bd)
        assert np.array_equal(d.download(), o.data), k
        n += 1
    assert n == 193


def test_forward_reference_vectors(r, ctx):
    """tests/golden/fwd_tx_r01.npz: the CUDA forward transforms against coefficients made by the reference's own
    1-D functions (tests/golden/make_fwd_tx_golden.py), all 193 pairs"""
    g = dict(np.load(os.path.join(HERE, "golden", "fwd_tx_r01.npz")))
    n = 0
    for k in sorted(g):
        if not k.startswith("x_"):
            continue
        s, t, bd = (int(v) for v in k.split("_")[1:])
        assert np.array_equal(r.forward_transform(ctx, g[k], s, t, bd), g["y" + k[1:]]), k
        n += 1
    assert n == 193


def test_inverse_add_reference_vectors(r, ctx, oracle, golden):
    """inv_txfm2d_add cases of the reference with both clamps biting"""
    for k in sorted(golden):
        if not k.startswith("ac_"):
            continue
        s, t, bd = (int(v) for v in k.split("_")[1:])
        before, after = golden["d" + k[2:]], golden["o" + k[2:]]
        h, w = before.shape
        _, d = twin_plane(r, oracle, ctx, w, h, before.dtype.itemsize, bd, data=before, xpad=0, ypad=0, pad=False)
        r.inverse_transform_add(ctx, golden[k], d, np.array([(0, 0)], dtype=r.ITX_JOB), s, t, bd)
        c = d.cfg
        got = d.download()[c.yorigin:c.yorigin + h, c.xorigin:c.xorigin + w]
        assert np.array_equal(got, after), k


@pytest.mark.parametrize("pb,bd", BD)
def test_inverse_all_pairs_match_oracle(r, ctx, oracle, pb, bd):
    """random, sparse and saturating coefficient blocks on random grid cells (also in the padding), batch
    sizes that do not fill the last CTA; whole allocation compared, so stray writes show"""
    rng = np.random.default_rng(50 + bd)
    L = oracle.lib()
    pairs = 0
    for s in range(19):
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        coded = min(w, 32) * min(h, 32)
        o, d = twin_plane(r, oracle, ctx, 256, 192, pb, bd, rng, xpad=64, ypad=64)
        for t in range(16):
            if not oracle.inv_txfm_valid(s, t):
                with pytest.raises(r.Rav1eCudaError) as e:
                    r.inverse_transform_add(ctx, np.zeros((1, coded), np.int32), d, np.zeros(1, dtype=r.ITX_JOB), s, t, bd)
                assert e.value.code == r.capi.R1C_ERR_UNIMPLEMENTED
                continue
            pairs += 1
            jobs = _grid_jobs(r, rng, int(rng.integers(1, 70)), w, h, 256 + 128, 192 + 128, -64, -64)
            n = jobs.size
            c = rng.integers(-(1 << (bd + 3)), 1 << (bd + 3), (n, coded))
            c = (c * rng.random((n, coded)) ** 2).astype(np.int64)
            for i in range(n):
                if i % 3 == 0:
                    c[i, int(rng.integers(1, coded + 1)):] = 0
            c[-1] = rng.choice([-(1 << (bd + 7)), (1 << (bd + 7)) - 1], coded)  # at the row clamp
            if n > 1:
                c[-2] = 0
                c[-2, 0] = (1 << (bd + 5))
            c = c.astype(np.int32)
            xy = np.stack([jobs["x"] + o.cfg.xorigin, jobs["y"] + o.cfg.yorigin], 1)
            oracle.inv_txfm_add_batch(c, o.data, xy, s, t, bd)
            r.inverse_transform_add(ctx, c, d, jobs, s, t, bd)
            assert np.array_equal(d.download(), o.data), (r.capi.TX_SIZES[s], r.capi.TX_TYPES[t], bd)
    assert pairs == 193


@pytest.mark.parametrize("pb,bd", BD)
def test_inverse_around_the_clamp_shortcut(r, ctx, oracle, pb, bd):
    """the kernels skip the clamps of a 1-D transform whose inputs have an L1 norm below range - 128 (itx1d.cuh): rows
    and columns whose norm sits just below, at and above that bound, many per call so that the lanes of a warp disagree,
    for every 1-D type and length; whole allocation against the oracle"""
    rng = np.random.default_rng(70 + bd)
    L = oracle.lib()
    hi_row, hi_col = (1 << (bd + 7)) - 1, (1 << (max(bd + 6, 16) - 1)) - 1
    for s, t in ((0, 0), (1, 0), (1, 3), (2, 0), (2, 3), (3, 0), (4, 0), (7, 1), (8, 2), (9, 0), (16, 12), (12, 0), (1, 9), (2, 10)):
        if not oracle.inv_txfm_valid(s, t):
            continue
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        cw, ch = min(w, 32), min(h, 32)
        o, d = twin_plane(r, oracle, ctx, 512, 256, pb, bd, rng, xpad=16, ypad=16)
        jobs = _grid_jobs(r, rng, 96, w, h, 512, 256)
        n = jobs.size
        c = np.zeros((n, ch, cw), dtype=np.int64)
        for i in range(n):
            kind = i % 4
            if kind == 0:  # one row (the DC row or a random one) with a norm around the row bound, spread over a few coefficients
                row = 0 if i % 8 == 0 else int(rng.integers(0, ch))
                total = hi_row - 128 + int(rng.integers(-300, 301))
                k = int(rng.integers(1, min(cw, 6) + 1))
                parts = rng.multinomial(total, np.ones(k) / k)
                c[i, row, rng.permutation(cw)[:k]] = parts * rng.choice([-1, 1], k)
            elif kind == 1:  # a DC that drives one column's inputs to the column bound after the row pass
                c[i, 0, 0] = int(rng.integers(hi_col // 2, hi_col * 8)) * int(rng.choice([-1, 1]))
            elif kind == 2:  # every row near the bound: the column norms are far above theirs
                for row in range(ch):
                    c[i, row, int(rng.integers(0, cw))] = (hi_row - 128 + int(rng.integers(-64, 65))) * int(rng.choice([-1, 1]))
            else:  # ordinary residual
                c[i] = rng.integers(-300, 301, (ch, cw)) * (rng.random((ch, cw)) < 0.2)
        c = c.reshape(n, -1).astype(np.int32)
        xy = np.stack([jobs["x"] + o.cfg.xorigin, jobs["y"] + o.cfg.yorigin], 1)
        oracle.inv_txfm_add_batch(c, o.data, xy, s, t, bd)
        r.inverse_transform_add(ctx, c, d, jobs, s, t, bd)
        assert np.array_equal(d.download(), o.data), (s, t)


@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10)])
def test_dequantize_fused_into_inverse(r, ctx, oracle, pb, bd):
    """r1c_dequant_inv_txfm_add_batch == dequantize (src/quantize.rs:318-334) then inverse_transform_add, for
    every log_tx_scale (0: up to 16x16 ..., 1: 32x32 / 16x64 ..., 2: 64x64 ...) and negative levels"""
    rng = np.random.default_rng(90 + bd)
    L = oracle.lib()
    for s, t, dcq, acq in ((0, 3, 8, 9), (1, 0, 40, 52), (2, 1, 300, 367), (3, 0, 96, 120), (4, 0, 1336, 1828),
                           (9, 0, 24, 30), (17, 0, 60, 71), (16, 11, 16, 20), (13, 9, 500, 600)):
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        coded = min(w, 32) * min(h, 32)
        o, d = twin_plane(r, oracle, ctx, 256, 192, pb, bd, rng, xpad=0, ypad=0)
        jobs = _grid_jobs(r, rng, 23, w, h, 256, 192)
        n = jobs.size
        q = (rng.integers(-40, 41, (n, coded)) * (rng.random((n, coded)) < 0.2)).astype(np.int32)
        q[:, 0] = rng.integers(-60, 61, n)
        lts = r.capi.load().r1c_log_tx_scale(s)
        rc = np.stack([oracle.dequantize(q[i], dcq, acq, lts) for i in range(n)])
        xy = np.stack([jobs["x"] + o.cfg.xorigin, jobs["y"] + o.cfg.yorigin], 1)
        oracle.inv_txfm_add_batch(rc, o.data, xy, s, t, bd)
        r.inverse_transform_add(ctx, q, d, jobs, s, t, bd, dequant=(dcq, acq))
        assert np.array_equal(d.download(), o.data), (s, t)
    with pytest.raises(r.Rav1eCudaError):
        r.inverse_transform_add(ctx, q, d, jobs, s, t, bd, dequant=(0, 5))


def test_inverse_errors_and_empty(r, ctx, oracle):
    rng = np.random.default_rng(9)
    o, d = twin_plane(r, oracle, ctx, 64, 64, 1, 8, rng, xpad=8, ypad=8)
    r.inverse_transform_add(ctx, np.zeros((0, 64), np.int32), d, np.zeros(0, dtype=r.ITX_JOB), 1, 0, 8)
    assert np.array_equal(d.download(), o.data)
    far = o.cfg.stride  # past the right edge of the allocation
    with pytest.raises(r.Rav1eCudaError) as e:
        r.inverse_transform_add(ctx, np.zeros((1, 64), np.int32), d, np.array([(far, 0)], dtype=r.ITX_JOB), 1, 0, 8)
    assert e.value.code == r.capi.R1C_ERR_BOUNDS
    with pytest.raises(r.Rav1eCudaError):  # 8-bit plane, 10-bit call
        r.inverse_transform_add(ctx, np.zeros((1, 64), np.int32), d, np.zeros(1, dtype=r.ITX_JOB), 1, 0, 10)
    with pytest.raises(r.Rav1eCudaError):
        r.inverse_transform_add(ctx, np.zeros((1, 64), np.int32), d, np.zeros(1, dtype=r.ITX_JOB), 19, 0, 8)
    assert np.array_equal(d.download(), o.data)


# (tx_size, tx_type, tolerance) rows of the reference's `roundtrips`, src/transform/mod.rs:665-708
ROUNDTRIPS = ([(0, t, 0) for t in (0, 1, 2, 3, 9, 10, 11, 12, 13)]
              + [(1, 0, 1), (1, 1, 1), (1, 2, 1), (1, 3, 1), (1, 9, 0), (1, 10, 0), (1, 11, 0), (1, 12, 0), (1, 13, 1)]
              + [(2, 0, 1), (2, 1, 1), (2, 2, 1), (2, 3, 1), (2, 9, 0), (2, 10, 1), (2, 11, 1)]
              + [(3, 0, 2), (3, 9, 0)]
              + [(5, 0, 1), (6, 0, 1), (13, 0, 1), (14, 0, 1), (7, 0, 1), (8, 0, 1), (15, 0, 2), (16, 0, 2),
                 (9, 0, 2), (10, 0, 2)])


def test_roundtrips_on_device(r, ctx, oracle):
    """diff + forward transform of (source - prediction), then inverse + add onto the prediction, gives the
    source back within the reference's tolerance"""
    rng = np.random.default_rng(77)
    L = oracle.lib()
    o_src, d_src = twin_plane(r, oracle, ctx, 256, 128, 1, 8, rng, xpad=0, ypad=0)
    for s, t, tol in ROUNDTRIPS:
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        o_dst, d_dst = twin_plane(r, oracle, ctx, 256, 128, 1, 8, rng, xpad=0, ypad=0)
        cells = _grid_jobs(r, rng, 40, w, h, 256, 128)
        tj = np.zeros(cells.size, dtype=r.TX_JOB)
        tj["in_x"], tj["in_y"], tj["pred_x"], tj["pred_y"] = cells["x"], cells["y"], cells["x"], cells["y"]
        freq = r.diff_forward_transform(ctx, d_src, d_dst, tj, s, t, 8)
        r.inverse_transform_add(ctx, freq.reshape(cells.size, -1), d_dst, cells, s, t, 8)
        c = d_dst.cfg
        got = d_dst.download()[c.yorigin:c.yorigin + 128, c.xorigin:c.xorigin + 256].astype(np.int64)
        for x, y in zip(cells["x"], cells["y"]):
            err = np.abs(got[y:y + h, x:x + w] - o_src.visible[y:y + h, x:x + w].astype(np.int64)).max()
            assert err <= tol, (s, t, int(err))
