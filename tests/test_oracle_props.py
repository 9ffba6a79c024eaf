"""Independent cross-checks of the UNPINNED parts of the oracle (MC, forward
TX, ME): properties that must hold for any correct restatement, because the
reference has no golden vectors for them."""
import ctypes as C

import numpy as np
import pytest


def _dct2(n):
    k = np.arange(n)[:, None]
    i = np.arange(n)[None, :]
    m = np.cos(np.pi * (2 * i + 1) * k / (2 * n)) * np.sqrt(2 / n)
    m[0] /= np.sqrt(2)
    return m


def _dst4(n):
    k = np.arange(n)[:, None]
    i = np.arange(n)[None, :]
    return np.sin(np.pi * (2 * i + 1) * (2 * k + 1) / (4 * n)) * np.sqrt(2 / n)


def _dst7(n):
    k = np.arange(n)[:, None]
    i = np.arange(n)[None, :]
    return np.sin(np.pi * (2 * k + 1) * (i + 1) / (2 * n + 1)) * 2 / np.sqrt(2 * n + 1)


def _mat(t1d, n):
    if t1d == 0:
        return _dct2(n)
    if t1d == 1:
        return _dst7(4) if n == 4 else _dst4(n)
    return np.eye(n)


@pytest.mark.parametrize("t1d,n", [(0, 4), (0, 8), (0, 16), (0, 32), (0, 64), (1, 4), (1, 8), (1, 16)])
def test_txfm1d_is_orthonormal_transform(oracle, t1d, n):
    """The Daala kernels of forward.rs approximate the ORTHONORMAL DCT-II /
    DST-IV (DST-VII for n=4) to within integer rounding."""
    rng = np.random.default_rng(n * 7 + t1d)
    M = _mat(t1d, n)
    X = rng.integers(-30000, 30000, size=(300, n)).astype(np.int32)
    Y = np.stack([oracle.txfm1d(t1d, x) for x in X]).astype(np.float64)
    assert np.abs(Y - X @ M.T).max() < 12.0
    # linearity at small amplitude: T(0) == 0
    assert not oracle.txfm1d(t1d, np.zeros(n, np.int32)).any()


def test_txfm1d_unimplemented(oracle):
    for t1d, n in [(1, 32), (1, 64), (2, 64)]:
        with pytest.raises(ValueError):
            oracle.txfm1d(t1d, np.zeros(n, np.int32))
    x = np.arange(32, dtype=np.int32)
    assert np.array_equal(oracle.txfm1d(2, x), x)  # identity == copy, forward.rs:1686-1700


def test_fht_valid_pair_count(oracle):
    L = oracle.lib()
    n = sum(L.r1o_fht_valid(s, t) for s in range(19) for t in range(16))
    assert n == 193


# (col1d, colflip, row1d, rowflip) per TxType, src/transform/mod.rs:196-213
TYPES = [(0, 0, 0, 0), (1, 0, 0, 0), (0, 0, 1, 0), (1, 0, 1, 0), (1, 1, 0, 0), (0, 0, 1, 1),
         (1, 1, 1, 1), (1, 0, 1, 1), (1, 1, 1, 0), (2, 0, 2, 0), (0, 0, 2, 0), (2, 0, 0, 0),
         (1, 0, 2, 0), (2, 0, 1, 0), (1, 1, 2, 0), (2, 0, 1, 1)]
SHIFT_SUM = {  # total left-shift of the three stages per shift table row (forward.rs:22-40)
}


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_fht_matches_float_separable(oracle, bd):
    """2-D fht == row_tx( col_tx(x << s0) >> s1 ) << / >> s2 up to rounding, for
    every valid (size, type)."""
    L = oracle.lib()
    rng = np.random.default_rng(bd)
    A = [[3, 0, 0], [2, 0, 1], [0, 0, 3]]
    B = [[4, -1, 0], [2, 0, 1], [0, 0, 3]]
    Cc = [[4, -2, 0], [2, 0, 0], [0, 0, 2]]
    D = [[4, -1, -2], [2, 0, -1], [0, 0, 1]]
    table = {0: A, 3: Cc, 9: Cc, 10: Cc, 17: Cc, 18: Cc, 4: D, 11: D, 12: D}
    amp = (1 << bd) - 1
    for s in range(19):
        W, H = L.r1o_tx_width(s), L.r1o_tx_height(s)
        sh = table.get(s, B)[(bd - 8) // 2]
        for t in range(16):
            if not L.r1o_fht_valid(s, t):
                continue
            ct, cf, rt, rf = TYPES[t]
            x = rng.integers(-amp, amp + 1, size=(H, W)).astype(np.int16)
            y = oracle.fht(x, s, t, bd).astype(np.float64)
            f = x.astype(np.float64) * 2.0 ** sh[0]
            if cf:
                f = f[::-1, :]
            f = _mat(ct, H) @ f
            f = f * 2.0 ** sh[1]
            if rf:
                f = f[:, ::-1]
            f = f @ _mat(rt, W).T
            f = f * 2.0 ** sh[2]
            tol = 8.0 * max(1.0, 2.0 ** sh[2]) + 4
            assert np.abs(y - f).max() < tol, (s, t, bd, np.abs(y - f).max())


# ---------------------------------------------------------------- MC ---------
def _rand_plane(oracle, rng, w, h, pb, bd):
    p = oracle.OPlane(w, h, 0, 0, 16, 16, pb)
    p.data[...] = rng.integers(0, 1 << bd, size=p.data.shape).astype(p.dtype)
    return p


def _put(oracle, src, x, y, w, h, cf, rf, mx, my, bd):
    dst = np.zeros((h, w), dtype=src.dtype)
    oracle.lib().r1o_put_8tap(dst.ctypes.data, w, src.ptr(x, y), src.cfg.stride, w, h, cf, rf, mx,
                              my, bd, src.px_bytes)
    return dst


def _prep(oracle, src, x, y, w, h, cf, rf, mx, my, bd):
    tmp = np.zeros((h, w), dtype=np.int16)
    oracle.lib().r1o_prep_8tap(tmp.ctypes.data, src.ptr(x, y), src.cfg.stride, w, h, cf, rf, mx, my,
                               bd, src.px_bytes)
    return tmp


@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10), (2, 12)])
def test_mc_identities(oracle, pb, bd):
    rng = np.random.default_rng(bd)
    src = _rand_plane(oracle, rng, 96, 80, pb, bd)
    ib = 2 if bd == 12 else 4
    for (w, h) in [(4, 4), (8, 16), (32, 8), (64, 64)]:
        blk = src.visible[10:10 + h, 20:20 + w]
        # frac (0,0): put == copy, prep == px << ib   (mc.rs:224-229, 320-327)
        assert np.array_equal(_put(oracle, src, 20, 10, w, h, 0, 0, 0, 0, bd), blk)
        assert np.array_equal(_prep(oracle, src, 20, 10, w, h, 0, 0, 2, 1, bd),
                              (blk.astype(np.int32) << ib).astype(np.int16))
        # mc_avg(prep, prep) at frac 0 reproduces the block
        t = _prep(oracle, src, 20, 10, w, h, 0, 0, 0, 0, bd)
        d = np.zeros((h, w), dtype=src.dtype)
        oracle.lib().r1o_mc_avg(d.ctypes.data, w, t.ctypes.data, t.ctypes.data, w, h, bd, pb)
        assert np.array_equal(d, blk)
        # bilinear half-pel horizontal == rounded mean of neighbours
        a = src.visible[10:10 + h, 20:20 + w].astype(np.int32)
        b = src.visible[10:10 + h, 21:21 + w].astype(np.int32)
        want = (a + b + 1) >> 1
        got = _put(oracle, src, 20, 10, w, h, 8, 0, 3, 3, bd).astype(np.int32)
        # two-stage rounding (7-ib then ib) of 64a+64b: may differ from the
        # single-stage mean by at most the double-rounding LSB
        assert np.abs(got - want).max() <= 1
        # vertical bilinear half-pel: single rounding by 7 -> exact mean
        b = src.visible[11:11 + h, 20:20 + w].astype(np.int32)
        got = _put(oracle, src, 20, 10, w, h, 0, 8, 3, 3, bd).astype(np.int32)
        assert np.array_equal(got, (a + b + 1) >> 1)


def test_mc_spec_rounding_constant_plane(oracle):
    """On a constant plane every filter (taps sum to 128) must return the
    constant for put, and const << ib for prep, in all four frac cases."""
    for pb, bd in [(1, 8), (2, 10), (2, 12)]:
        src = oracle.OPlane(64, 64, 0, 0, 16, 16, pb, fill=(1 << bd) - 3)
        ib = 2 if bd == 12 else 4
        for cf, rf in [(0, 5), (7, 0), (3, 11), (15, 15)]:
            for mx in range(4):
                for w, h in [(4, 8), (16, 4), (16, 16)]:
                    p = _put(oracle, src, 8, 8, w, h, cf, rf, mx, (mx + 1) % 4, bd)
                    assert (p == (1 << bd) - 3).all()
                    t = _prep(oracle, src, 8, 8, w, h, cf, rf, mx, (mx + 1) % 4, bd)
                    assert (t == ((1 << bd) - 3) << ib).all()


def test_mv_params(oracle):
    """src/predict.rs:267-284: floor shift for the integer part, 1/16 fracs,
    clamp applied to (origin - 3) then +3."""
    cfg = oracle.OPlane(64, 48, 0, 0, 16, 16, 1).cfg
    rf, cf, sx, sy = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
    f = oracle.lib().r1o_mv_params
    f(C.byref(cfg), 8, 8, -9, 13, C.byref(rf), C.byref(cf), C.byref(sx), C.byref(sy))
    assert (rf.value, cf.value) == (((-9) - (-2 << 3)) << 1, (13 - 8) << 1)
    assert (sx.value, sy.value) == (8 + 1, 8 - 2)
    f(C.byref(cfg), 8, 8, -8000, 8000, C.byref(rf), C.byref(cf), C.byref(sx), C.byref(sy))
    assert (sx.value, sy.value) == (cfg.width + 3, -cfg.yorigin + 3)
    # chroma plane (xdec=ydec=1): shift 4, frac already in 1/16
    ccfg = oracle.OPlane(32, 24, 1, 1, 8, 8, 1).cfg
    f(C.byref(ccfg), 4, 4, 21, -5, C.byref(rf), C.byref(cf), C.byref(sx), C.byref(sy))
    assert (rf.value, cf.value, sx.value, sy.value) == (5, 11, 4 - 1, 4 + 1)


# ---------------------------------------------------------------- ME ---------
def test_mv_rate(oracle):
    Mv, f = oracle.Mv, oracle.lib().r1o_mv_rate
    assert f(Mv(0, 0), Mv(0, 0), 0) == 0
    assert f(Mv(1, 0), Mv(0, 0), 0) == 0          # 1 >> 1 == 0
    assert f(Mv(-1, 0), Mv(0, 0), 0) == 2         # -1 >> 1 == -1 -> 1 bit
    assert f(Mv(8, -8), Mv(0, 0), 0) == 2 * 3 + 2 * 3
    assert f(Mv(8, -8), Mv(0, 0), 1) == 2 * 4 + 2 * 4


def test_full_search_finds_shift_and_first_min(oracle):
    rng = np.random.default_rng(3)
    ref = oracle.OPlane(128, 96, 0, 0, 88, 88, 1)
    ref.visible[...] = rng.integers(0, 256, size=(96, 128))
    ref.pad()
    org = oracle.OPlane(128, 96, 0, 0, 88, 88, 1)
    org.visible[...] = np.roll(ref.visible, (-3, 5), axis=(0, 1))  # org(y,x) = ref(y+3, x-5)
    org.pad()
    best, cost = oracle.Mv(0, 0), C.c_uint64(2 ** 64 - 1)
    pmv = (oracle.Mv * 2)()
    rc = oracle.lib().r1o_full_search(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(),
                                      C.byref(ref.cfg), 32 - 16, 32 + 16, 32 - 16, 32 + 16, 16, 16,
                                      32, 32, 1, 0, C.byref(pmv), 0, 1, C.byref(best), C.byref(cost))
    assert rc == 0 and (best.row, best.col, cost.value) == (3 * 8, -5 * 8, 0)
    # flat planes: every candidate ties -> the first raster position wins
    flat = oracle.OPlane(128, 96, 0, 0, 88, 88, 1)
    best, cost = oracle.Mv(0, 0), C.c_uint64(2 ** 64 - 1)
    oracle.lib().r1o_full_search(flat.origin_ptr(), C.byref(flat.cfg), flat.origin_ptr(),
                                 C.byref(flat.cfg), 30, 40, 28, 36, 8, 8, 32, 32, 2, 0,
                                 C.byref(pmv), 0, 1, C.byref(best), C.byref(cost))
    assert (best.row, best.col, cost.value) == ((28 - 32) * 8, (30 - 32) * 8, 0)
    # a window leaving the allocation is an error (PlaneRegion::new asserts)
    rc = oracle.lib().r1o_full_search(flat.origin_ptr(), C.byref(flat.cfg), flat.origin_ptr(),
                                      C.byref(flat.cfg), -200, 40, 28, 36, 8, 8, 32, 32, 1, 0,
                                      C.byref(pmv), 0, 1, C.byref(best), C.byref(cost))
    assert rc == -1


# ------------------------------------------------------------- RDO distortion -----
def _np_bias(imp, w_in_imp_b, w_in_b, h_in_b, bo_x, bo_y, wmi, hmi):
    """compute_distortion_bias restated independently with numpy float32 scalars (src/rdo.rs:477-508)"""
    total = np.float32(0)
    for y in range(bo_y, min(bo_y + hmi, h_in_b)):
        for x in range(bo_x, min(bo_x + wmi, w_in_b)):
            total = np.float32(total + imp[y // 2 * w_in_imp_b + x // 2])
    mean = np.float32(total / np.float32(wmi * hmi))
    return float(np.float32(mean / np.float32(3))) + 0.65


@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10), (2, 12)])
def test_sse_wxh_and_cdef_dist_against_numpy(oracle, pb, bd):
    rng = np.random.default_rng(900 + bd)
    W, H = 96, 64
    a = oracle.OPlane(W, H, 0, 0, 16, 16, pb)
    b = oracle.OPlane(W, H, 0, 0, 16, 16, pb)
    a.data[...] = rng.integers(0, 1 << bd, a.data.shape)
    b.data[...] = np.clip(a.data.astype(np.int64) + rng.integers(-40, 41, a.data.shape), 0, (1 << bd) - 1)
    w_in_b, h_in_b = W // 4, H // 4
    w_imp = w_in_b // 2
    imp = (rng.random(w_imp * (h_in_b // 2)) * 4).astype(np.float32)
    for (x, y, w, h) in [(0, 0, 4, 4), (8, 4, 4, 16), (16, 8, 16, 8), (32, 0, 64, 64), (88, 56, 8, 8), (4, 4, 8, 8)]:
        va, vb = a.visible[y:y + h, x:x + w].astype(np.int64), b.visible[y:y + h, x:x + w].astype(np.int64)
        assert oracle.sse_wxh(a, x, y, b, x, y, w, h) == int(((va - vb) ** 2).sum())
        bw, bh = min(w, 8), min(h, 8)
        want = 0
        for by in range(h // bh):
            for bx in range(w // bw):
                d = va[by * bh:(by + 1) * bh, bx * bw:(bx + 1) * bw] - vb[by * bh:(by + 1) * bh, bx * bw:(bx + 1) * bw]
                bias = _np_bias(imp, w_imp, w_in_b, h_in_b, (x + bx * bw) // 4, (y + by * bh) // 4, bw // 4, bh // 4)
                want += int(float(int((d * d).sum())) * bias)
        assert oracle.sse_wxh(a, x, y, b, x, y, w, h, imp, w_imp, w_in_b, h_in_b) == want
        if w % 8 == 0 and h % 8 == 0 and x % 8 == 0 and y % 8 == 0:
            cs = bd - 8
            want = 0
            for by in range(h // 8):
                for bx in range(w // 8):
                    s = va[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8]
                    d = vb[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8]
                    ss, sd = int(s.sum()), int(d.sum())
                    s2, d2, sdx = int((s * s).sum()), int((d * d).sum()), int((s * d).sum())
                    svar = float(s2 - ((ss * ss + 32) >> 6))
                    dvar = float(d2 - ((sd * sd + 32) >> 6))
                    boost = (4033.0 / 16384.0) * (svar + dvar + float(16384 << (2 * cs))) / np.sqrt(
                        float(16265089 << (4 * cs)) + svar * dvar)
                    value = int(float(d2 + s2 - 2 * sdx) * float(boost) + 0.5)
                    bias = _np_bias(imp, w_imp, w_in_b, h_in_b, (x + bx * 8) // 4, (y + by * 8) // 4, 2, 2)
                    want += int(float(value) * bias)
            assert oracle.cdef_dist_wxh(a, x, y, b, x, y, w, h, bd, imp, w_imp, w_in_b, h_in_b) == want
            assert oracle.cdef_dist_wxh(a, x, y, a, x, y, w, h, bd) == 0


def test_sse_wxh_chroma_uses_decimated_importance_blocks(oracle):
    """4:2:0 chroma: an importance block is 4x4 chroma px and covers 2x2 MI of the frame (src/rdo.rs:296-301)"""
    rng = np.random.default_rng(77)
    a = oracle.OPlane(32, 32, 1, 1, 8, 8, 1)
    b = oracle.OPlane(32, 32, 1, 1, 8, 8, 1)
    a.data[...] = rng.integers(0, 256, a.data.shape)
    b.data[...] = rng.integers(0, 256, b.data.shape)
    w_in_b = h_in_b = 16
    imp = (rng.random(8 * 8) * 2).astype(np.float32)
    x, y, w, h = 8, 4, 8, 8  # chroma px
    want = 0
    for by in range(2):
        for bx in range(2):
            d = (a.visible[y + by * 4:y + by * 4 + 4, x + bx * 4:x + bx * 4 + 4].astype(np.int64) -
                 b.visible[y + by * 4:y + by * 4 + 4, x + bx * 4:x + bx * 4 + 4])
            bias = _np_bias(imp, 8, w_in_b, h_in_b, (x + bx * 4) // 2, (y + by * 4) // 2, 2, 2)
            want += int(float(int((d * d).sum())) * bias)
    assert oracle.sse_wxh(a, x, y, b, x, y, w, h, imp, 8, w_in_b, h_in_b) == want


def test_tx_dist_scaling_and_coded_area(oracle):
    rng = np.random.default_rng(5)
    L = oracle.lib()
    for ts in range(19):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        c = rng.integers(-20000, 20000, w * h)
        rc = rng.integers(-20000, 20000, coded)
        bits = 2 * (3 - ((w * h > 256) + (w * h > 1024)))
        want = (int(((c[:coded] - rc) ** 2).sum()) + (1 << (bits - 1))) >> bits
        assert oracle.tx_dist(c, rc, ts) == want
    # the i32 product wraps, then sign-extends into the u64 sum (src/encoder.rs:1215-1217)
    c = np.array([70000] + [0] * 15)
    rc = np.array([-70000] + [0] * 15)
    wrapped = ((140000 * 140000 + 2 ** 31) % 2 ** 32) - 2 ** 31
    assert oracle.tx_dist(c, rc, 0) == ((wrapped % 2 ** 64) + 32) >> 6


def test_diamond_search_walks_to_the_shift(oracle):
    """diamond_me_search (src/me.rs:693-785): on a smooth plane shifted by (+16, -24) px the full-pel diamond
    started at (0, 0) ends on the true shift; the sub-pel stage then stays there (cost 0 with lambda 0)"""
    rng = np.random.default_rng(3)
    W, H = 192, 128
    yy, xx = np.mgrid[0:H, 0:W]
    v = np.rint(127 + 60 * np.sin(2 * np.pi * xx / 300) + 60 * np.sin(2 * np.pi * yy / 260) + rng.integers(0, 2, size=(H, W)))
    ref = oracle.OPlane(W, H, 0, 0, 88, 88, 1)
    ref.visible[...] = v.astype(np.uint8)
    ref.pad()
    org = oracle.OPlane(W, H, 0, 0, 88, 88, 1)
    org.visible[...] = np.roll(ref.visible, (-16, 24), axis=(0, 1))  # org(y, x) = ref(y + 16, x - 24)
    org.pad()
    mv_range = (-8 * 64, 8 * 64, -8 * 64, 8 * 64)
    pmv = [(0, 0), (0, 0)]
    mv, cost, iters = oracle.diamond_search(org, ref, (64, 48), 32, 32, [(0, 0)], pmv, 0, False, False, mv_range, False, 8)
    start = oracle.diamond_search(org, ref, (64, 48), 32, 32, [(0, 0)], pmv, 0, False, False, (0, 0, 0, 0), False, 8)[1]
    # a 4-neighbour descent may stop one pixel off on a diagonal valley
    assert abs(mv[0] - 16 * 8) <= 8 and abs(mv[1] + 24 * 8) <= 8 and cost < start // 10 and iters > 3
    mv2, cost2, _ = oracle.diamond_search(org, ref, (64, 48), 32, 32, [mv, (16 * 8, -24 * 8)], pmv, 0, True, False, mv_range,
                                          True, 8)
    assert mv2 == (16 * 8, -24 * 8) and cost2 == 0  # get_best_predictor picks the exact shift, the diamond keeps it
    # every predictor outside the range: the reference would assert; the restatement returns u64::MAX
    mv3, cost3, _ = oracle.diamond_search(org, ref, (64, 48), 32, 32, [(9999, 0)], pmv, 0, False, False, (-8, 8, -8, 8), False, 8)
    assert cost3 >= (1 << 64) - 1 or mv3 != (9999, 0)


def test_subset_predictors_and_save_block_motion(oracle):
    """get_subset_predictors (src/me.rs:82-174): zero MV first, coarse MVs quantised toward zero, zero neighbours dropped,
    median of the available neighbours; save_block_motion (src/encoder.rs:1432-1444) clips at the field edge"""
    f = np.zeros((4, 6, 2), dtype=np.int16)
    oracle.save_block_motion(f, (4, 2), 16, 16, (-13, 27))  # 4x4 MI block hanging over both edges
    assert (f[2:, 4:] == (-13, 27)).all() and (f[:2] == 0).all() and (f[:, :4] == 0).all()
    prev = np.zeros((4, 6, 2), dtype=np.int16)
    prev[1, 1] = (5, 5)
    p = oracle.get_subset_predictors((0, 0), [(-9, 15)], f, prev)
    assert p == [(0, 0), (-8, 8)]  # corner: no neighbours, prev right / bottom / co-located are zero
    f[1, 0] = (16, -8)
    f[0, 1] = (0, 0)
    f[0, 2] = (-24, 40)
    p = oracle.get_subset_predictors((1, 1), [], f, prev)
    # left (16,-8), top zero (dropped but in the median), top-right (-24,40); median = (-8/3, 32/3) -> (-2, 10) -> (0, 8);
    # then the co-located previous MV
    assert p == [(0, 0), (16, -8), (-24, 40), (0, 8), (5, 5)]


def test_importance_propagation_properties(oracle):
    """compute_block_importances (src/api/internal.rs:823-1075): with zero MVs and ref == frame the inter cost is 0,
    so every block hands intra_cost + future importance to the same cell; a MV of (+4, +4) px splits it 1/4 each
    over four cells; shares that fall outside the grid are dropped; finalize is log2(1 + imp / intra)"""
    w, h = 12, 9
    rng = np.random.default_rng(3)
    pl = oracle.OPlane(w * 8, h * 8, 0, 0, 88, 88, 1)
    pl.visible[...] = rng.integers(0, 256, size=(h * 8, w * 8))
    pl.pad()
    intra = rng.integers(1, 5000, size=(h, w)).astype(np.uint32)
    fut = rng.random((h, w)).astype(np.float32) * 100
    mvs = np.zeros((2 * h, 2 * w, 2), dtype=np.int16)
    out = oracle.importance_propagate(pl, pl, mvs, intra, fut, 1, 8, np.zeros((h, w), np.float32))
    assert np.array_equal(out, intra.astype(np.float32) + fut)
    out2 = oracle.importance_propagate(pl, pl, mvs, intra, fut, 2, 8, out.copy())
    assert np.array_equal(out2, out + (intra.astype(np.float32) + fut) / np.float32(2))
    # constant plane: any MV keeps the inter cost at 0; (+32, +32) in 1/8 px = half a block each way
    flat = oracle.OPlane(w * 8, h * 8, 0, 0, 88, 88, 1)
    flat.data[...] = 77
    mvs[...] = 32
    got = oracle.importance_propagate(flat, flat, mvs, intra, np.zeros((h, w), np.float32), 1, 8, np.zeros((h, w), np.float32))
    want = np.zeros((h + 1, w + 1), np.float64)
    a = intra.astype(np.float64) / 4
    for dy in (0, 1):
        for dx in (0, 1):
            want[dy:dy + h, dx:dx + w] += a
    assert np.allclose(got, want[:h, :w], rtol=1e-6)
    imp = got.copy()
    oracle.importance_finalize(intra, imp)
    assert np.allclose(imp, np.log2(1 + got / intra.astype(np.float32)), rtol=1e-6)
    z = np.array([[5.0]], np.float32)
    assert oracle.importance_finalize(np.zeros((1, 1), np.uint32), z)[0, 0] == 0.0


def test_satd_i16_lanes_equal_reference_semantics_until_they_wrap(oracle):
    """r1o_satd_i16 (the fork's generated kernels, i16 lanes for every pixel type, build/kernel/gen/dist.rs:171-350):
    the merged last butterfly (|a + b| + |a - b| = 2 max(|a|, |b|)) makes it EQUAL to get_satd_ref whenever no lane
    leaves 16 bits -- always for 8-bit input (|term| <= 64 * 255) -- and different for 10- / 12-bit blocks of high
    amplitude (a flat 10-bit difference of 1023: the DC halves are 32 * 1023 each and their doubling wraps)"""
    rng = np.random.default_rng(23)
    L = oracle.lib()
    for pb, bd in ((1, 8), (2, 10), (2, 12)):
        a = oracle.OPlane(64, 64, 0, 0, 8, 8, pb)
        b = oracle.OPlane(64, 64, 0, 0, 8, 8, pb)
        top = (1 << bd) - 1
        yy, xx = np.indices((64, 64))
        board = ((xx + yy) & 1) * top
        for name, x, y in (("noise", None, None), ("random", rng.integers(0, top + 1, size=(64, 64)), rng.integers(0, top + 1, size=(64, 64))),
                           ("flat", np.full((64, 64), top), np.zeros((64, 64), int)), ("checkerboard", board, top - board)):
            if x is None:
                x = rng.integers(0, top + 1, size=(64, 64))
                y = np.clip(x + rng.integers(-60, 61, size=(64, 64)), 0, top)
            a.visible[...] = x
            b.visible[...] = y
            differ = 0
            for bs in range(22):
                w, h = L.r1o_block_width(bs), L.r1o_block_height(bs)
                if w > 64 or h > 64:
                    continue
                differ += oracle.satd(a, 0, 0, b, 0, 0, w, h) != oracle.satd_i16(a, 0, 0, b, 0, 0, w, h)
            if bd == 8 or name == "noise":
                assert differ == 0, (bd, name)
            if bd >= 10 and name in ("flat", "checkerboard"):
                assert differ > 0, (bd, name)
