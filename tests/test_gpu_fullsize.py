"""GPU parity at BASELINE.json's FULL sizes (configs[1..3]): the whole 1080p / 4K frames go through the CUDA path;
the oracle (OpenMP, all host threads) checks whole frames where it finishes in seconds (MC, forward TX, quantize,
SAD / SATD) and a stratified sample of blocks where it does not (full search +-64), and size-independent properties
cover the rest: a globally shifted frame is found at cost 0 by every block, superblock-row shards reproduce the whole
frame, a frac (0, 0) prediction is a copy, identity transforms are linear."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import bench_kernels
from gpu_util import twin_plane

pytestmark = pytest.mark.gpu
ALL = sum(1 << b for b in range(22) if b not in (13, 14, 15))


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


def _layout(oracle, W, H):
    L = oracle.lib()
    off, cols, rows = (C.c_int64 * 22)(), (C.c_int32 * 22)(), (C.c_int32 * 22)()
    total = L.r1o_frame_me_layout(W, H, ALL, C.byref(off), C.byref(cols), C.byref(rows))
    return total, [(b, off[b], cols[b], rows[b]) for b in range(22) if off[b] >= 0]


@pytest.mark.parametrize("pb,bd,W,H,frac", [(1, 8, 1920, 1080, 0.004), (2, 10, 3840, 2160, 0.0006), (2, 12, 1920, 1080, 0.002)])
def test_frame_me_full_size_sampled_against_oracle(r, ctx, oracle, pb, bd, W, H, frac):
    """configs[1] (and its 4K 10-bit counterpart): full search +-64, 19 sizes, SATD sub-pel tail on the whole frame;
    a stratified sample of blocks of every size -- incl. the frame's corners and edges -- against the oracle"""
    cur, ref = bench_kernels.synth_frame_pair(1, W, H, bd)
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=cur)
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, data=ref)
    lam = 160 << (bd - 8)
    p = r.make_frame_me_params(W, H, 64, 64, lam)
    got = r.frame_me(ctx, d_org, d_ref, p, bd)
    total, layout = _layout(oracle, W, H)
    assert got.size == total
    rng = np.random.default_rng(9)
    idx = []
    for b, off, cols, rows in layout:
        n = cols * rows
        k = max(6, int(round(n * frac)))
        pick = set(int(v) for v in rng.choice(n, size=min(n, k), replace=False))
        pick |= {0, cols - 1, n - cols, n - 1, (rows // 2) * cols, (rows // 2) * cols + cols - 1}  # corners, edge middles
        idx += [off + v for v in sorted(pick)]
    idx = np.array(sorted(set(idx)), dtype=np.int64)
    mv, cost = oracle.frame_me(o_org, o_ref, 64, 64, lam, ALL, 1, 1, 0, bd, idx)
    assert np.array_equal(got["cost"][idx], cost[idx])
    assert np.array_equal(got["mv"]["row"][idx], mv[idx, 0]) and np.array_equal(got["mv"]["col"][idx], mv[idx, 1])
    # superblock-row shards reproduce the whole frame, and the run is deterministic
    sb_rows = (H + 63) // 64
    parts = np.zeros(total, dtype=got.dtype)
    parts["cost"] = np.iinfo(np.uint64).max
    for lo, hi in ((0, sb_rows // 3), (sb_rows // 3, sb_rows)):
        ps = r.make_frame_me_params(W, H, 64, 64, lam, sb_rows=(lo, hi))
        assert r.frame_me(ctx, d_org, d_ref, ps, bd, results=parts) is parts
    assert hashlib.sha256(parts.tobytes()).digest() == hashlib.sha256(got.tobytes()).digest()


def test_frame_me_full_size_finds_a_global_shift(r, ctx, oracle):
    """1080p: ref = cur moved by (+21, -37) px; with lambda = 0 every block of every size whose window holds the
    shifted block must come back with the shift and cost 0 (full-pel only)"""
    W, H, dx, dy = 1920, 1080, 21, -37
    rng = np.random.default_rng(12)
    big = rng.integers(0, 256, size=(H + 128, W + 128), dtype=np.uint8)
    cur = big[64:64 + H, 64:64 + W]
    ref = big[64 - dy:64 - dy + H, 64 - dx:64 - dx + W]  # ref(y + dy, x + dx) = cur(y, x)
    _, d_org = twin_plane(r, oracle, ctx, W, H, 1, 8, data=cur)
    _, d_ref = twin_plane(r, oracle, ctx, W, H, 1, 8, data=ref)
    p = r.make_frame_me_params(W, H, 64, 64, 0, refine=0)
    got = r.frame_me(ctx, d_org, d_ref, p, 8)
    L = oracle.lib()
    _, layout = _layout(oracle, W, H)
    for b, off, cols, rows in layout:
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        g = got[off:off + cols * rows].reshape(rows, cols)
        ys, xs = np.meshgrid(np.arange(rows) * bh, np.arange(cols) * bw, indexing="ij")
        inside = (xs + dx >= 0) & (xs + dx + bw <= W) & (ys + dy >= 0) & (ys + dy + bh <= H)
        assert inside.sum() > 0.9 * inside.size
        assert (g["cost"][inside] == 0).all(), (bw, bh)
        assert (g["mv"]["col"][inside] == 8 * dx).all() and (g["mv"]["row"][inside] == 8 * dy).all(), (bw, bh)


def test_mc_full_4k_10bit_frame_against_oracle(r, ctx, oracle):
    """configs[2]: the 4K 10-bit luma plane tiled by 16x16 and by 64x32 blocks, random fracs / filter pairs / +-32 px
    offsets: put_8tap of the whole frame, prep_8tap x2 + mc_avg of the whole frame, against the oracle (OpenMP);
    and frac (0, 0) is a copy of the displaced source"""
    W, H, bd, pb = 3840, 2160, 10, 2
    rng = np.random.default_rng(33)
    o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, rng)
    L = oracle.lib()
    for (w, h) in ((16, 16), (64, 32)):
        xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
        n = xs.size
        jobs = np.zeros(n, dtype=r.MC_JOB)
        jobs["dst_x"], jobs["dst_y"] = xs.ravel(), ys.ravel()
        jobs["src_x"] = jobs["dst_x"] + rng.integers(-32, 33, size=n)
        jobs["src_y"] = jobs["dst_y"] + rng.integers(-32, 33, size=n)
        jobs["w"], jobs["h"] = w, h
        jobs["col_frac"], jobs["row_frac"] = rng.integers(0, 16, size=n), rng.integers(0, 16, size=n)
        jobs["mode_x"], jobs["mode_y"] = rng.integers(0, 4, size=n), rng.integers(0, 4, size=n)
        jobs["tmp_off"] = np.arange(n) * (w * h)
        oj = np.stack([jobs[k].astype(np.int32) for k in ("dst_x", "dst_y", "src_x", "src_y", "w", "h", "col_frac",
                                                          "row_frac", "mode_x", "mode_y")], axis=1).copy()
        toff = jobs["tmp_off"].astype(np.int64)
        o_dst, d_dst = twin_plane(r, oracle, ctx, W, H, pb, bd, pad=False)
        r.put_8tap_batch(ctx, d_dst, d_src, jobs, bd)
        L.r1o_mc_batch(0, o_dst.origin_ptr(), C.byref(o_dst.cfg), None, None, o_src.origin_ptr(), C.byref(o_src.cfg),
                       oj.ctypes.data, n, bd, pb)
        assert np.array_equal(d_dst.download(), o_dst.data), ("put", w, h)
        got1 = np.zeros(n * w * h, np.int16)
        r.prep_8tap_batch(ctx, got1, d_src, jobs, bd)
        want1 = np.zeros(n * w * h, np.int16)
        L.r1o_mc_batch(1, None, None, want1.ctypes.data, toff.ctypes.data, o_src.origin_ptr(), C.byref(o_src.cfg),
                       oj.ctypes.data, n, bd, pb)
        assert np.array_equal(got1, want1), ("prep", w, h)
        # compound: second prediction = the same blocks from another displacement, then mc_avg
        jobs2 = jobs.copy()
        jobs2["src_x"] = jobs["dst_x"] + rng.integers(-32, 33, size=n)
        jobs2["col_frac"] = rng.integers(0, 16, size=n)
        got2 = np.zeros(n * w * h, np.int16)
        r.prep_8tap_batch(ctx, got2, d_src, jobs2, bd)
        aj = np.zeros(n, dtype=r.AVG_JOB)
        aj["dst_x"], aj["dst_y"], aj["w"], aj["h"], aj["tmp_off"] = jobs["dst_x"], jobs["dst_y"], w, h, jobs["tmp_off"]
        r.mc_avg_batch(ctx, d_dst, got1, got2, aj, bd)
        oa = np.stack([aj[k].astype(np.int32) for k in ("dst_x", "dst_y", "w", "h")], axis=1).copy()
        L.r1o_avg_batch(o_dst.origin_ptr(), C.byref(o_dst.cfg), got1.ctypes.data, got2.ctypes.data, toff.ctypes.data,
                        oa.ctypes.data, n, bd, pb)
        assert np.array_equal(d_dst.download(), o_dst.data), ("avg", w, h)
    # frac (0, 0): a copy of the source displaced by (-5, +9) for the whole frame (64x64 jobs)
    xs, ys = np.meshgrid(np.arange(0, W, 64), np.arange(0, H - 63, 64))
    jobs = np.zeros(xs.size, dtype=r.MC_JOB)
    jobs["dst_x"], jobs["dst_y"], jobs["src_x"], jobs["src_y"] = xs.ravel(), ys.ravel(), xs.ravel() - 5, ys.ravel() + 9
    jobs["w"], jobs["h"] = 64, 64
    o_dst, d_dst = twin_plane(r, oracle, ctx, W, H, pb, bd, pad=False)
    r.put_8tap_batch(ctx, d_dst, d_src, jobs, bd)
    hh = (H // 64) * 64
    yo, xo = int(o_src.cfg.yorigin), int(o_src.cfg.xorigin)
    moved = o_src.data[yo + 9:yo + 9 + hh, xo - 5:xo - 5 + W]
    got = d_dst.download()[int(o_dst.cfg.yorigin):int(o_dst.cfg.yorigin) + hh, int(o_dst.cfg.xorigin):int(o_dst.cfg.xorigin) + W]
    assert np.array_equal(got, moved)


def test_fwd_txfm_quant_full_4k_10bit_residual(r, ctx, oracle):
    """configs[3]: the 4:2:0 4K 10-bit residual (12.4 M samples) through forward_transform for three transform sizes /
    types, then quantize + dequantize of the coefficients, all against the oracle; IDTX is linear"""
    from quant_util import zigzag_scan
    n_samples = 3840 * 2160 * 3 // 2
    rng = np.random.default_rng(44)
    res = rng.integers(-1023, 1024, size=n_samples).astype(np.int16)
    L = oracle.lib()
    for ts, tt in ((1, 0), (2, 3), (12, 0)):  # 8x8 DCT_DCT, 16x16 ADST_ADST, 64x32 DCT_DCT
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        n = n_samples // (w * h)
        blk = res[: n * w * h].reshape(n, h, w)
        co = r.forward_transform(ctx, blk, ts, tt, 10)
        want = oracle.fht_batch(blk, ts, tt, 10)
        assert np.array_equal(co.reshape(-1), want.reshape(-1)), (ts, tt)
        scan = zigzag_scan(min(w, 32), min(h, 32))
        qc = r.QuantizationContext().update(ctx.lib, ts, 0, 120, 147)
        q = qc.quantize(ctx, co.reshape(n, -1), scan)
        oq, orc = oracle.quantize_batch(oracle.quant_ctx(ts, 0, 120, 147), ts, scan, co.reshape(n, -1), want_rcoeffs=True)
        assert np.array_equal(q, oq), ts
        assert np.array_equal(r.dequantize(ctx, q, ts, 120, 147), orc), ts
    # identity transform: fht(a + b) == fht(a) + fht(b) exactly (pure shifts; 8-bit 4x4 shifts are all left shifts)
    a = rng.integers(-255, 256, size=(n_samples // 16, 4, 4)).astype(np.int16)
    b = rng.integers(-255, 256, size=a.shape).astype(np.int16)
    fa, fb, fab = (r.forward_transform(ctx, x, 0, 9, 8) for x in (a, b, (a + b).astype(np.int16)))
    assert np.array_equal(fab, fa + fb)


def test_reconstruction_chain_full_4k_10bit_plane(r, ctx, oracle):
    """the encoder's per-block chain over a whole 4K 10-bit luma plane: diff + forward transform, quantize,
    dequantize, inverse transform added onto the prediction (src/encoder.rs:1143-1209).  The reconstructed plane is
    compared with the oracle's chain bit for bit, and -- independently of the oracle -- must come back to the source
    within the quantizer's reach (the prediction is off by up to 24; mean |error| < step / 16, max <= step / 2)."""
    from quant_util import zigzag_scan
    W, H, bd, pb = 3840, 2160, 10, 2
    rng = np.random.default_rng(66)
    o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    L = oracle.lib()
    for ts, tt, dcq, acq in ((1, 0, 40, 52), (3, 0, 96, 120), (12, 0, 24, 30), (7, 2, 16, 20)):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        noise = rng.integers(-24, 25, size=(H, W))
        pred = np.clip(o_src.visible.astype(np.int64) + noise, 0, 1023).astype(np.uint16)
        o_rec, d_rec = twin_plane(r, oracle, ctx, W, H, pb, bd, data=pred)
        xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
        n = xs.size
        tj = np.zeros(n, dtype=r.TX_JOB)
        tj["in_x"] = tj["pred_x"] = xs.ravel()
        tj["in_y"] = tj["pred_y"] = ys.ravel()
        cells = np.zeros(n, dtype=r.ITX_JOB)
        cells["x"], cells["y"] = xs.ravel(), ys.ravel()
        coded = min(w, 32) * min(h, 32)
        scan = zigzag_scan(min(w, 32), min(h, 32))
        # device chain
        co = r.diff_forward_transform(ctx, d_src, d_rec, tj, ts, tt, bd)
        q = r.QuantizationContext().update(ctx.lib, ts, 0, dcq, acq).quantize(ctx, co.reshape(n, -1), scan)
        rc = r.dequantize(ctx, q, ts, dcq, acq)
        r.inverse_transform_add(ctx, rc, d_rec, cells, ts, tt, bd)
        got = d_rec.download()
        # oracle chain
        res = np.zeros((n, h, w), np.int16)
        src_v, pred_v = o_src.visible.astype(np.int32), pred.astype(np.int32)
        for i in range(n):
            x, y = int(cells["x"][i]), int(cells["y"][i])
            res[i] = src_v[y:y + h, x:x + w] - pred_v[y:y + h, x:x + w]
        oco = oracle.fht_batch(res, ts, tt, bd).reshape(n, -1)
        assert np.array_equal(co.reshape(n, -1), oco), ts
        oq, orc = oracle.quantize_batch(oracle.quant_ctx(ts, 0, dcq, acq), ts, scan, oco, want_rcoeffs=True)
        assert np.array_equal(q, oq) and np.array_equal(rc, orc), ts
        assert orc.shape[1] == coded
        xy = np.stack([cells["x"] + o_rec.cfg.xorigin, cells["y"] + o_rec.cfg.yorigin], 1)
        oracle.inv_txfm_add_batch(orc, o_rec.data, xy, ts, tt, bd)
        assert np.array_equal(got, o_rec.data), ts
        # the domain's own check: inside the tiled area the reconstruction is close to the source.  Not for
        # 64-point sizes: the fork's fht writes them dense with stride W while quantize / inverse read them 32 wide
        # (reproduced on purpose), so their reconstruction is not meaningful -- parity above still holds
        if max(w, h) == 64:
            continue
        c = d_rec.cfg
        vis = got[c.yorigin:c.yorigin + (H // h) * h, c.xorigin:c.xorigin + (W // w) * w].astype(np.int64)
        err = np.abs(vis - src_v[:(H // h) * h, :(W // w) * w])
        assert err.mean() < acq / 16 and err.max() <= acq / 2, (ts, float(err.mean()), int(err.max()))


def test_sad_satd_full_4k_10bit_frame_against_oracle(r, ctx, oracle):
    """every 8x8 and 32x16 block of a 4K 10-bit frame pair, reference displaced by +-32 px: SAD and SATD of the whole
    frame against the oracle, and sum of the 8x8 SADs of an undisplaced tiling == SAD of the frame (a checksum of
    checksums computed independently in numpy)"""
    W, H, bd, pb = 3840, 2160, 10, 2
    rng = np.random.default_rng(55)
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, rng)
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng)
    L = oracle.lib()
    for b in (3, 8):
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        xs, ys = np.meshgrid(np.arange(0, W, bw), np.arange(0, H - bh + 1, bh))
        n = xs.size
        jobs = np.zeros(n, dtype=r.DIST_JOB)
        jobs["org_x"], jobs["org_y"], jobs["bsize"] = xs.ravel(), ys.ravel(), b
        jobs["ref_x"] = jobs["org_x"] + rng.integers(-32, 33, size=n)
        jobs["ref_y"] = jobs["org_y"] + rng.integers(-32, 33, size=n)
        oj = np.stack([jobs[k].astype(np.int32) for k in ("org_x", "org_y", "ref_x", "ref_y", "bsize")], axis=1).copy()
        for satd, fn in ((0, r.sad_batch), (1, r.satd_batch)):
            want = np.zeros(n, dtype=np.uint32)
            L.r1o_dist_batch(satd, o_org.origin_ptr(), C.byref(o_org.cfg), o_ref.origin_ptr(), C.byref(o_ref.cfg),
                             oj.ctypes.data, n, pb, want.ctypes.data)
            assert np.array_equal(fn(ctx, d_org, d_ref, jobs, bd), want), (b, satd)
    xs, ys = np.meshgrid(np.arange(0, W, 8), np.arange(0, H, 8))
    jobs = np.zeros(xs.size, dtype=r.DIST_JOB)
    jobs["org_x"], jobs["org_y"], jobs["ref_x"], jobs["ref_y"], jobs["bsize"] = xs.ravel(), ys.ravel(), xs.ravel(), ys.ravel(), 3
    total = int(r.sad_batch(ctx, d_org, d_ref, jobs, bd).astype(np.uint64).sum())
    assert total == int(np.abs(o_org.visible.astype(np.int64) - o_ref.visible.astype(np.int64)).sum())
