"""The CPU oracle against known answers made by executing the reference's own Rust text
(tests/golden/{mc,me,quant,rdo}_r02.npz, written by tests/golden/make_ref_golden.py from /root/reference/src/mc.rs,
me.rs, predict.rs, quantize.rs, rdo.rs through tests/rustlite.py).  These files pin the oracle for MC, motion search,
quantize / dequantize and RDO distortion to the reference; tests/test_gpu_ref_golden.py checks the CUDA path against the
same files on the GPU box.  Bit-exact everywhere."""
import ctypes as C
import os

import numpy as np
import pytest

from golden_util import G, load_golden, oplane

U64_MAX = (1 << 64) - 1


# ------------------------------------------------------------------------------------------------- MC ------
@pytest.mark.parametrize("bd", [8, 10, 12])
def test_oracle_put_prep_avg_match_reference(oracle, bd):
    g = load_golden("mc")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    src = oplane(oracle, g["src_%d" % bd], pb, xpad, ypad)
    L = oracle.lib()
    off = 0
    for sx, sy, w, h, cf, rf, mx, my in g["cases_%d" % bd]:
        w, h = int(w), int(h)
        dst = np.zeros((h, w), dtype=src.dtype)
        L.r1o_put_8tap(dst.ctypes.data, w, src.ptr(int(sx), int(sy)), src.cfg.stride, w, h, int(cf), int(rf), int(mx), int(my), bd, pb)
        assert np.array_equal(dst.reshape(-1), g["put_%d" % bd][off:off + w * h]), ("put", sx, sy, w, h, cf, rf, mx, my)
        tmp = np.zeros(w * h, dtype=np.int16)
        L.r1o_prep_8tap(tmp.ctypes.data, src.ptr(int(sx), int(sy)), src.cfg.stride, w, h, int(cf), int(rf), int(mx), int(my), bd, pb)
        assert np.array_equal(tmp, g["prep_%d" % bd][off:off + w * h]), ("prep", sx, sy, w, h, cf, rf, mx, my)
        off += w * h
    assert off == g["put_%d" % bd].size
    off = 0
    for w, h in g["avg_sizes_%d" % bd]:
        n = int(w) * int(h)
        t1 = np.ascontiguousarray(g["avg_t1_%d" % bd][off:off + n])
        t2 = np.ascontiguousarray(g["avg_t2_%d" % bd][off:off + n])
        dst = np.zeros(n, dtype=src.dtype)
        L.r1o_mc_avg(dst.ctypes.data, int(w), t1.ctypes.data, t2.ctypes.data, int(w), int(h), bd, pb)
        assert np.array_equal(dst, g["avg_out_%d" % bd][off:off + n]), ("avg", w, h)
        off += n


def _oracle_predict(oracle, plane, po, w, h, mv, bd, pb):
    """r1o_mv_params (predict_inter's get_params + clamp) followed by put_8tap, REGULAR / REGULAR"""
    L = oracle.lib()
    rf, cf, sx, sy = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
    L.r1o_mv_params(C.byref(plane.cfg), po[0], po[1], mv[0], mv[1], C.byref(rf), C.byref(cf), C.byref(sx), C.byref(sy))
    dst = np.zeros((h, w), dtype=plane.dtype)
    L.r1o_put_8tap(dst.ctypes.data, w, plane.ptr(sx.value, sy.value), plane.cfg.stride, w, h, cf.value, rf.value, 0, 0, bd, pb)
    return dst.reshape(-1)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_oracle_predict_inter_params_and_clamp_match_reference(oracle, bd):
    g = load_golden("mc")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    src = oplane(oracle, g["src_%d" % bd], pb, xpad, ypad)
    csrc = oplane(oracle, g["csrc_%d" % bd], pb, 24, 24, xdec=1, ydec=1)
    for plane, ck, ok in ((src, "pi_cases_%d", "pi_out_%d"), (csrc, "pc_cases_%d", "pc_out_%d")):
        off = 0
        for po_x, po_y, w, h, mr, mc in g[ck % bd]:
            n = int(w) * int(h)
            got = _oracle_predict(oracle, plane, (int(po_x), int(po_y)), int(w), int(h), (int(mr), int(mc)), bd, pb)
            assert np.array_equal(got, g[ok % bd][off:off + n]), (ck, po_x, po_y, w, h, mr, mc)
            off += n


# ------------------------------------------------------------------------------------------------- ME ------
def test_oracle_mv_range_and_rate_match_reference(oracle):
    g = load_golden("me")
    L = oracle.lib()
    for c, want in zip(g["range_cases"], g["range_out"]):
        out = (C.c_int64 * 4)()
        L.r1o_mv_range(*(int(v) for v in c), C.byref(out))
        assert tuple(out) == tuple(int(v) for v in want), tuple(c)
    for (ar, ac, br, bc, hp), want in zip(g["rate_cases"], g["rate_out"]):
        assert L.r1o_mv_rate(oracle.Mv(int(ar), int(ac)), oracle.Mv(int(br), int(bc)), int(hp)) == int(want)


def _fs(oracle, org, ref, j, pb):
    L = oracle.lib()
    b = int(j[2])
    bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
    best, cost = oracle.Mv(0, 0), C.c_uint64(U64_MAX)
    pmv = (oracle.Mv * 2)(oracle.Mv(int(j[9]), int(j[10])), oracle.Mv(int(j[11]), int(j[12])))
    rc = L.r1o_full_search(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(), C.byref(ref.cfg), int(j[3]), int(j[4]),
                           int(j[5]), int(j[6]), bw, bh, int(j[0]), int(j[1]), int(j[7]), int(j[8]), C.byref(pmv),
                           int(j[13]), pb, C.byref(best), C.byref(cost))
    assert rc == 0
    return (best.row, best.col), cost.value


@pytest.mark.parametrize("bd", [8, 10])
def test_oracle_full_search_matches_reference(oracle, bd):
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    org = oplane(oracle, g["org_%d" % bd], pb, xpad, ypad)
    ref = oplane(oracle, g["ref_%d" % bd], pb, xpad, ypad)
    for j, mv, cost in zip(g["fs_jobs_%d" % bd], g["fs_mv_%d" % bd], g["fs_cost_%d" % bd]):
        assert _fs(oracle, org, ref, j, pb) == ((int(mv[0]), int(mv[1])), int(cost)), tuple(j)


def test_oracle_full_search_ties_match_reference(oracle):
    g = load_golden("me")
    flat = oplane(oracle, np.full((32, 32), 77, dtype=np.uint16), 1, 40, 40)
    for j, mv, cost in zip(g["tie_jobs"], g["tie_mv"], g["tie_cost"]):
        assert _fs(oracle, flat, flat, j, 1) == ((int(mv[0]), int(mv[1])), int(cost)), tuple(j)


def _subpel_cost(oracle, org, ref, j, mr, cand, bd, pb, use_satd=None):
    L = oracle.lib()
    b = int(j[2])
    bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
    pmv = (oracle.Mv * 2)(oracle.Mv(int(j[4]), int(j[5])), oracle.Mv(int(j[6]), int(j[7])))
    rng = (C.c_int64 * 4)(*(int(v) for v in mr))
    return L.r1o_subpel_cost(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(), C.byref(ref.cfg), int(j[0]), int(j[1]),
                             bw, bh, oracle.Mv(int(cand[0]), int(cand[1])), C.byref(pmv), int(j[3]),
                             int(j[8]) if use_satd is None else use_satd, int(j[9]), C.byref(rng), bd, pb)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_oracle_subpel_cost_and_telescopic_match_reference(oracle, bd):
    """get_mv_rd_cost through predict_inter (src/me.rs:787-856) and telescopic_subpel_search (:858-941)"""
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    org = oplane(oracle, g["org_%d" % bd], pb, xpad, ypad)
    ref = oplane(oracle, g["ref_%d" % bd], pb, xpad, ypad)
    jobs, ranges = g["sp_jobs_%d" % bd], g["sp_range_%d" % bd]
    for k, j in enumerate(jobs):
        for c, want in zip(g["sp_cand_%d" % bd][k], g["sp_cost_%d" % bd][k]):
            assert _subpel_cost(oracle, org, ref, j, ranges[k], c, bd, pb) == int(want), (k, tuple(c))
        # telescopic: steps 8, 4, 2 (, 1) x 8 neighbours in (i, j) raster order, strict `<`
        best, lowest = (int(j[10]), int(j[11])), int(g["tele_start_cost_%d" % bd][k])
        for step in [8, 4, 2] + ([1] if int(j[9]) else []):
            centre = best
            for i in range(3):
                for jj in range(3):
                    if i == 1 and jj == 1:
                        continue
                    cand = (centre[0] + step * (i - 1), centre[1] + step * (jj - 1))
                    c = _subpel_cost(oracle, org, ref, j, ranges[k], cand, bd, pb)
                    if c < lowest:
                        lowest, best = c, cand
        assert (best, lowest) == (tuple(int(v) for v in g["tele_mv_%d" % bd][k]), int(g["tele_cost_%d" % bd][k])), k


@pytest.mark.parametrize("bd", [8, 10])
def test_oracle_diamond_search_matches_reference(oracle, bd):
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    org = oplane(oracle, g["org_%d" % bd], pb, xpad, ypad)
    ref = oplane(oracle, g["ref_%d" % bd], pb, xpad, ypad)
    L = oracle.lib()
    for k, j in enumerate(g["dia_jobs_%d" % bd]):
        b = int(j[2])
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        preds = [(int(p[0]), int(p[1])) for p in g["dia_pred_%d" % bd][k]]
        mv, cost, _it = oracle.diamond_search(org, ref, (int(j[0]), int(j[1])), bw, bh, preds,
                                              [(int(j[4]), int(j[5])), (int(j[6]), int(j[7]))], int(j[3]), int(j[8]), int(j[9]),
                                              tuple(int(v) for v in j[10:14]), int(g["dia_sub_%d" % bd][k]), bd)
        assert (mv, cost) == (tuple(int(v) for v in g["dia_mv_%d" % bd][k]), int(g["dia_cost_%d" % bd][k])), k


# --------------------------------------------------------------------------------------------- quantize ----
def test_oracle_quantize_dequantize_match_reference(oracle):
    g = load_golden("quant")
    L = oracle.lib()
    for k, (ts, tt, qi, intra, bd, n) in enumerate(g["cases"]):
        ctx = [int(v) for v in g["ctx_%d" % k]]
        qc = oracle.quant_ctx(int(ts), int(intra), ctx[1], ctx[2])
        got = [qc.log_tx_scale, qc.dc_quant, qc.ac_quant, qc.dc_offset, qc.ac_offset_eob, qc.ac_offset0, qc.ac_offset1] + \
            list(qc.dc_mul_add) + list(qc.ac_mul_add)
        assert got == ctx, (k, "QuantizationContext::update")
        assert L.r1o_log_tx_scale(int(ts)) == ctx[0]
        q, r = oracle.quantize_batch(qc, int(ts), g["scan_%d" % k], g["coeffs_%d" % k], want_rcoeffs=True)
        assert np.array_equal(q, g["q_%d" % k]), (k, "quantize")
        assert np.array_equal(r, g["r_%d" % k]), (k, "dequantize")
        for b in range(int(n)):
            assert np.array_equal(oracle.dequantize(g["q_%d" % k][b], ctx[1], ctx[2], ctx[0]), g["r_%d" % k][b])
    for (x, d), want in zip(g["divu_cases"], g["divu_out"]):
        assert L.r1o_divu_pair(int(x), C.byref(oracle.divu_gen(int(d)))) == int(want), (x, d)


# -------------------------------------------------------------------------------------------------- RDO ----
@pytest.mark.parametrize("bd", [8, 10, 12])
def test_oracle_rdo_distortion_matches_reference(oracle, bd):
    g = load_golden("rdo")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    a = oplane(oracle, g["a_%d" % bd], pb, xpad, ypad)
    b = oplane(oracle, g["b_%d" % bd], pb, xpad, ypad)
    imp = g["imp_%d" % bd]
    for (x, y, w, h, kind, wb), want in zip(g["cases_%d" % bd], g["out_%d" % bd]):
        args = (imp, W // 8, W // 4, H // 4) if wb else ()
        if kind == 0:
            got = oracle.sse_wxh(a, int(x), int(y), b, int(x), int(y), int(w), int(h), *args)
        else:
            got = oracle.cdef_dist_wxh(a, int(x), int(y), b, int(x), int(y), int(w), int(h), bd, *args)
        assert got == int(want), (x, y, w, h, kind, wb)
    ca = oplane(oracle, g["a_%d" % bd][::2, ::2], pb, 8, 8, xdec=1, ydec=1)
    cb = oplane(oracle, g["b_%d" % bd][::2, ::2], pb, 8, 8, xdec=1, ydec=1)
    for (x, y, w, h, _kind, wb), want in zip(g["ccases_%d" % bd], g["cout_%d" % bd]):
        args = (imp, W // 8, W // 4, H // 4) if wb else ()
        assert oracle.sse_wxh(ca, int(x), int(y), cb, int(x), int(y), int(w), int(h), *args) == int(want), (x, y, w, h)


def test_golden_files_are_present():
    for name in G:
        assert os.path.isfile(G[name]), name
