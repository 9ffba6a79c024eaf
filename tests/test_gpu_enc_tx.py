"""GPU parity of the fused encode_tx_block kernel (csrc/enc_tx.cu*, reference src/encoder.rs:1143-1224) through the
C ABI: quantized coefficients, tx-domain distortion and the reconstructed plane against the oracle's chain
(diff -> fht -> quantize -> dequantize -> inverse add -> tx_dist), for every fused (TxSize, TxType) pair, three bit
depths and several scan orders, and against the chain of the separate CUDA entry points at full plane size."""
import numpy as np
import pytest

from gpu_util import twin_plane
from quant_util import col_scan, random_scan, row_scan, zigzag_scan

pytestmark = pytest.mark.gpu
BD = [(1, 8), (2, 10), (2, 12)]
FUSED_SIZES = list(range(19))  # every TxSize; the five with a 64-point side go through enc_tx64.cuh


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


def _oracle_chain(oracle, o_src, o_rec, cells, ts, tt, bd, is_intra, dcq, acq, scan, reconstruct=True):
    """-> (qcoeffs, tx_dist); o_rec.data is updated in place when `reconstruct`"""
    L = oracle.lib()
    w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
    n = len(cells)
    res = np.zeros((n, h, w), np.int16)
    sv, pv = o_src.visible.astype(np.int32), o_rec.visible.astype(np.int32)
    for i, (x, y) in enumerate(cells):
        res[i] = sv[y:y + h, x:x + w] - pv[y:y + h, x:x + w]
    co = oracle.fht_batch(res, ts, tt, bd).reshape(n, -1)
    q, rc = oracle.quantize_batch(oracle.quant_ctx(ts, is_intra, dcq, acq), ts, scan, co, want_rcoeffs=True)
    dist = np.array([oracle.tx_dist(co[i], rc[i], ts) for i in range(n)], dtype=np.uint64)
    if reconstruct:
        xy = np.array(cells, dtype=np.int64).reshape(-1, 2) + [o_rec.cfg.xorigin, o_rec.cfg.yorigin]
        oracle.inv_txfm_add_batch(rc, o_rec.data, xy, ts, tt, bd)
    return q, dist


def _cells(rng, w, h, W, H, n=None):
    cells = [(x, y) for y in range(0, H - h + 1, h) for x in range(0, W - w + 1, w)]
    order = rng.permutation(len(cells))
    if n is not None:
        order = order[:n]
    return [cells[k] for k in order]


def _jobs(r, cells):
    tj = np.zeros(len(cells), dtype=r.TX_JOB)
    xy = np.array(cells, dtype=np.int32).reshape(-1, 2)
    tj["in_x"] = tj["pred_x"] = xy[:, 0]
    tj["in_y"] = tj["pred_y"] = xy[:, 1]
    return tj


def test_valid_pairs(r, ctx):
    """fused where forward and inverse both exist"""
    lib = ctx.lib
    n = 0
    for ts in range(19):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        for tt in range(16):
            want = int(lib.r1c_fwd_txfm_valid(ts, tt) and lib.r1c_inv_txfm_valid(ts, tt))
            assert lib.r1c_encode_tx_block_valid(ts, tt) == want, (ts, tt)
            n += want
    assert n > 100
    assert lib.r1c_encode_tx_block_valid(-1, 0) == 0 and lib.r1c_encode_tx_block_valid(0, 16) == 0


@pytest.mark.parametrize("pb,bd", BD)
def test_every_fused_pair_against_oracle_chain(r, ctx, oracle, pb, bd):
    """all fused (size, type) pairs: random cells of a 160x128 plane (a ragged last slab), prediction = source + noise,
    zig-zag scan; quantizers from fine to coarse"""
    rng = np.random.default_rng(100 + bd)
    W, H = 160, 128
    lib = ctx.lib
    o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    for ts in FUSED_SIZES:
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        scan = zigzag_scan(min(w, 32), min(h, 32))  # av1_get_coded_tx_size: the scan covers the coded area
        for tt in range(16):
            if not lib.r1c_encode_tx_block_valid(ts, tt):
                continue
            amp = int(rng.choice([3, 20, 200])) << (bd - 8)
            pred = np.clip(o_src.visible.astype(np.int64) + rng.integers(-amp, amp + 1, size=(H, W)), 0, (1 << bd) - 1)
            o_rec, d_rec = twin_plane(r, oracle, ctx, W, H, pb, bd, data=pred.astype(o_src.dtype))
            cells = _cells(rng, w, h, W, H, n=int(rng.integers(1, 3 * (W // w) * (H // h) // 4 + 2)))
            dcq, acq = (int(v) for v in rng.integers(4, 400, size=2))
            is_intra = int(rng.integers(0, 2))
            qc = r.QuantizationContext().update(lib, ts, is_intra, dcq, acq)
            q, dist = r.encode_tx_block_batch(ctx, d_src, d_rec, _jobs(r, cells), ts, tt, bd, qc, scan)
            oq, odist = _oracle_chain(oracle, o_src, o_rec, cells, ts, tt, bd, is_intra, dcq, acq, scan)
            assert np.array_equal(q, oq), (ts, tt, "qcoeffs")
            assert np.array_equal(dist, odist), (ts, tt, "tx_dist")
            assert np.array_equal(d_rec.download(), o_rec.data), (ts, tt, "rec")


def test_scan_orders_flags_and_extremes(r, ctx, oracle):
    """row / column / random scans; reconstruct = 0 leaves the plane alone; no tx_dist; flat and extreme residuals
    (saturated prediction against an empty source and the reverse) at 12 bits"""
    rng = np.random.default_rng(7)
    W, H, pb, bd = 128, 96, 2, 12
    lib = ctx.lib
    for ts, tt in ((0, 0), (1, 3), (2, 9), (3, 0), (7, 1), (14, 12), (15, 10), (9, 0), (4, 0), (11, 10), (12, 0), (17, 5), (18, 1)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        cw, ch = min(w, 32), min(h, 32)
        for case, scan in enumerate((row_scan(cw, ch), col_scan(cw, ch), random_scan(rng, cw * ch))):
            src = rng.integers(0, 4096, size=(H, W))
            pred = rng.integers(0, 4096, size=(H, W))
            if case == 0:
                src[:, : W // 2], pred[:, : W // 2] = 4095, 0
                src[:, W // 2:], pred[:, W // 2:] = 0, 4095
            elif case == 1:
                pred = src.copy()  # zero residual: all-zero coefficients, eob search finds nothing
                pred[: H // 2] = np.clip(src[: H // 2] + rng.integers(-1, 2, size=(H // 2, W)), 0, 4095)
            o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, data=src.astype(np.uint16))
            o_rec, d_rec = twin_plane(r, oracle, ctx, W, H, pb, bd, data=pred.astype(np.uint16))
            cells = _cells(rng, w, h, W, H)
            dcq, acq = (8, 9) if case == 1 else (int(v) for v in rng.integers(4, 2000, size=2))
            qc = r.QuantizationContext().update(lib, ts, 0, dcq, acq)
            before = d_rec.download().copy()
            q0, d0 = r.encode_tx_block_batch(ctx, d_src, d_rec, _jobs(r, cells), ts, tt, bd, qc, scan, reconstruct=False,
                                             want_tx_dist=False)
            assert d0 is None and np.array_equal(d_rec.download(), before), (ts, tt, "reconstruct=0 wrote the plane")
            q, dist = r.encode_tx_block_batch(ctx, d_src, d_rec, _jobs(r, cells), ts, tt, bd, qc, scan)
            oq, odist = _oracle_chain(oracle, o_src, o_rec, cells, ts, tt, bd, 0, dcq, acq, scan)
            assert np.array_equal(q0, oq) and np.array_equal(q, oq), (ts, tt, case, "qcoeffs")
            assert np.array_equal(dist, odist), (ts, tt, case, "tx_dist")
            assert np.array_equal(d_rec.download(), o_rec.data), (ts, tt, case, "rec")


def test_errors_and_empty(r, ctx, oracle):
    lib = ctx.lib
    _, d_a = twin_plane(r, oracle, ctx, 64, 64, 1, 8, np.random.default_rng(1))
    _, d_b = twin_plane(r, oracle, ctx, 64, 64, 1, 8, np.random.default_rng(2))
    _, d_c = twin_plane(r, oracle, ctx, 64, 64, 2, 10, np.random.default_rng(3))
    qc = r.QuantizationContext().update(lib, 1, 0, 40, 52)
    scan = zigzag_scan(8, 8)
    q, d = r.encode_tx_block_batch(ctx, d_a, d_b, np.zeros(0, dtype=r.TX_JOB), 1, 0, 8, qc, scan)
    assert q.shape == (0, 64) and d.shape == (0,)
    one = np.zeros(1, dtype=r.TX_JOB)
    with pytest.raises(r.Rav1eCudaError) as e:  # a 64-point side is a DCT
        qc64 = r.QuantizationContext().update(lib, 4, 0, 40, 52)
        r.encode_tx_block_batch(ctx, d_a, d_b, one, 4, 9, 8, qc64, zigzag_scan(32, 32))
    assert e.value.code == r.capi.R1C_ERR_UNIMPLEMENTED
    q64, _ = r.encode_tx_block_batch(ctx, d_a, d_b, one, 4, 0, 8, qc64, zigzag_scan(32, 32))  # one 64x64 block = the plane
    assert q64.shape == (1, 1024)
    with pytest.raises(r.Rav1eCudaError) as e:  # 32-point ADST does not exist
        qc32 = r.QuantizationContext().update(lib, 3, 0, 40, 52)
        r.encode_tx_block_batch(ctx, d_a, d_b, one, 3, 3, 8, qc32, zigzag_scan(32, 32))
    assert e.value.code == r.capi.R1C_ERR_UNIMPLEMENTED
    with pytest.raises(r.Rav1eCudaError) as e:  # pixel types differ
        r.encode_tx_block_batch(ctx, d_a, d_c, one, 1, 0, 8, qc, scan)
    assert e.value.code == r.capi.R1C_ERR_ARG
    with pytest.raises(r.Rav1eCudaError) as e:  # quantization context of another size class (log_tx_scale differs)
        r.encode_tx_block_batch(ctx, d_a, d_b, one, 3, 0, 8, qc, zigzag_scan(32, 32))
    assert e.value.code == r.capi.R1C_ERR_ARG
    bad = scan.copy()
    bad[5] = bad[6]
    with pytest.raises(r.Rav1eCudaError) as e:  # not a permutation
        r.encode_tx_block_batch(ctx, d_a, d_b, one, 1, 0, 8, qc, bad)
    assert e.value.code == r.capi.R1C_ERR_ARG
    far = np.zeros(1, dtype=r.TX_JOB)
    far["in_x"] = 64 + 88 + 512
    with pytest.raises(r.Rav1eCudaError) as e:
        r.encode_tx_block_batch(ctx, d_a, d_b, far, 1, 0, 8, qc, scan)
    assert e.value.code == r.capi.R1C_ERR_BOUNDS


def test_full_4k_10bit_plane_against_separate_kernels(r, ctx, oracle):
    """a whole 4K 10-bit luma plane tiled by 8x8 / 16x16 / 32x32 / 16x8 / 64x64 / 64x32 / 16x64: the fused call must leave the same qcoeffs and
    plane as r1c_diff_fwd_txfm -> r1c_quantize -> r1c_dequant_inv_txfm_add, and bring the plane back to the source
    within the quantizer's reach (the domain's own check, independent of both)"""
    W, H, bd, pb = 3840, 2160, 10, 2
    rng = np.random.default_rng(67)
    o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    lib = ctx.lib
    for ts, tt, dcq, acq in ((1, 0, 40, 52), (2, 3, 60, 70), (3, 0, 96, 120), (8, 11, 24, 30), (4, 0, 80, 100), (12, 0, 50, 60),
                             (17, 2, 30, 44)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        pred = np.clip(o_src.visible.astype(np.int64) + rng.integers(-24, 25, size=(H, W)), 0, 1023).astype(np.uint16)
        _, d_rec = twin_plane(r, oracle, ctx, W, H, pb, bd, data=pred)
        _, d_rec2 = twin_plane(r, oracle, ctx, W, H, pb, bd, data=pred)
        xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
        n = xs.size
        tj = np.zeros(n, dtype=r.TX_JOB)
        tj["in_x"] = tj["pred_x"] = xs.ravel()
        tj["in_y"] = tj["pred_y"] = ys.ravel()
        cells = np.zeros(n, dtype=r.ITX_JOB)
        cells["x"], cells["y"] = xs.ravel(), ys.ravel()
        scan = zigzag_scan(min(w, 32), min(h, 32))
        qc = r.QuantizationContext().update(lib, ts, 0, dcq, acq)
        q, dist = r.encode_tx_block_batch(ctx, d_src, d_rec, tj, ts, tt, bd, qc, scan)
        co = r.diff_forward_transform(ctx, d_src, d_rec2, tj, ts, tt, bd)
        q2 = qc.quantize(ctx, co.reshape(n, -1), scan)
        r.inverse_transform_add(ctx, q2, d_rec2, cells, ts, tt, bd, dequant=(dcq, acq))
        assert np.array_equal(q, q2), ts
        got = d_rec.download()
        assert np.array_equal(got, d_rec2.download()), ts
        rc = r.dequantize(ctx, q2, ts, dcq, acq)
        assert np.array_equal(dist, r.tx_distortion(ctx, co.reshape(n, -1), rc, ts)), ts
        if max(w, h) == 64:
            continue  # this fork's 64-point layouts do not reconstruct the source: parity only
        c = d_rec.cfg
        vis = got[c.yorigin:c.yorigin + (H // h) * h, c.xorigin:c.xorigin + (W // w) * w].astype(np.int64)
        err = np.abs(vis - o_src.visible[:(H // h) * h, :(W // w) * w].astype(np.int64))
        assert err.mean() < acq / 16 and err.max() <= acq / 2, (ts, float(err.mean()), int(err.max()))
