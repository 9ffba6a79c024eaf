"""Dev-container-only (skipped wherever /root/reference is absent, e.g. on the GPU box): the reference's own Rust text
for MC, motion search, quantize and RDO distortion is EXECUTED (tests/rustlite.py interprets src/mc.rs, me.rs,
predict.rs, quantize.rs, rdo.rs, dist.rs, frame/plane.rs, tiling/plane_region.rs ... read from /root/reference at test
time) and compared with (a) the CPU oracle on fresh random inputs and (b) the committed golden files
tests/golden/*_r02.npz, so that those files are known to be what the reference computes, not what the oracle does."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference tree not present")

from golden_util import load_golden, oplane  # noqa: E402

PIXEL = {8: "u8", 10: "u16", 12: "u16"}


def _X():
    import ref_rs
    return ref_rs


def _pair(X, oracle, bd, vis, xpad, ypad, xdec=0, ydec=0):
    I = X.interp(PIXEL[bd])
    h, w = vis.shape
    rp = X.RPlane(I, w, h, xdec, ydec, xpad, ypad)
    rp.set_visible(vis)
    rp.pad(w << xdec, h << ydec)
    op = oplane(oracle, vis, 1 if bd == 8 else 2, xpad, ypad, xdec, ydec)
    return I, rp, op


@pytest.mark.parametrize("bd", [8, 10])
def test_plane_new_pad_downsample_are_the_references(oracle, bd):
    """Plane::new geometry, pad, downsample_from (src/frame/plane.rs:215-314,399-423) run from their text"""
    X = _X()
    rng = np.random.default_rng(bd)
    for (w, h, xdec, ydec, xpad, ypad) in [(40, 24, 0, 0, 16, 8), (20, 12, 1, 1, 9, 5), (66, 30, 0, 0, 88, 88)]:
        vis = rng.integers(0, 1 << bd, size=(h, w))
        I, rp, op = _pair(X, oracle, bd, vis, xpad, ypad, xdec, ydec)
        for k in ("stride", "alloc_height", "width", "height", "xdec", "ydec", "xpad", "ypad", "xorigin", "yorigin"):
            assert rp.cfg(k) == getattr(op.cfg, k), k
        assert np.array_equal(rp.all(), op.data)
    vis = rng.integers(0, 1 << bd, size=(24, 40))
    I, rp, op = _pair(X, oracle, bd, vis, 16, 8)
    half = X.RPlane(I, 20, 12, 0, 0, 8, 4)
    I.call("Plane::downsample_from", half.s, rp.s)
    oh = oracle.OPlane(20, 12, 0, 0, 8, 4, op.px_bytes)
    oh.data[...] = 0
    assert oracle.lib().r1o_plane_downsample(oh.data.ctypes.data, C.byref(oh.cfg), op.data.ctypes.data, C.byref(op.cfg), op.px_bytes) == 0
    assert np.array_equal(half.all(), oh.data)
    assert I.calls["downsample_from"] >= 1 and I.calls["pad"] >= 1


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_sad_satd_text_vs_oracle(oracle, bd):
    """get_sad_ref / get_satd_ref (src/dist.rs:25-46,197-328) run from their text, all 22 block sizes"""
    X = _X()
    rng = np.random.default_rng(100 + bd)
    a = rng.integers(0, 1 << bd, size=(140, 140))
    b = np.where(rng.random((140, 140)) < 0.2, (1 << bd) - 1 - a, a + rng.integers(-9, 9, size=(140, 140))).clip(0, (1 << bd) - 1)
    I, ra, oa = _pair(X, oracle, bd, a, 8, 8)
    _, rb, ob = _pair(X, oracle, bd, b, 8, 8)
    L = oracle.lib()
    for bs in range(22):
        w, h = L.r1o_block_width(bs), L.r1o_block_height(bs)
        if w * h > 64 * 64 and bd != 8:
            continue  # keep the interpreted side in seconds
        assert X.get_sad(I, ra, (3, 2), rb, (5, 1), bs, bd) == oracle.sad(oa, 3, 2, ob, 5, 1, w, h), bs
        if w * h <= 64 * 32:
            assert X.get_satd(I, ra, (3, 2), rb, (5, 1), bs, bd) == oracle.satd(oa, 3, 2, ob, 5, 1, w, h), bs
    assert I.calls["get_satd_ref"] > 0 and I.calls["hadamard8_1d"] > 0


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_mc_text_vs_oracle_random(oracle, bd):
    """put_8tap_ref / prep_8tap_ref / mc_avg_ref (src/mc.rs:213-408) on fresh random cases"""
    X = _X()
    rng = np.random.default_rng(200 + bd)
    vis = rng.integers(0, 1 << bd, size=(40, 48))
    I, rp, op = _pair(X, oracle, bd, vis, 16, 16)
    L = oracle.lib()
    pb = op.px_bytes
    for k in range(24):
        w, h = [(2, 2), (4, 4), (8, 8), (4, 16), (16, 4), (10, 6), (2, 8), (16, 16)][k % 8]
        sx, sy = int(rng.integers(-6, 48 - w + 6)), int(rng.integers(-6, 40 - h + 6))
        cf, rf = int(rng.integers(0, 16)) * (k % 3 != 0), int(rng.integers(0, 16)) * (k % 4 != 1)
        mx, my = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        d = X.RPlane(I, w, h)
        X.put_8tap(I, d, (0, 0), rp, (sx, sy), w, h, cf, rf, mx, my, bd)
        want = np.zeros((h, w), dtype=op.dtype)
        L.r1o_put_8tap(want.ctypes.data, w, op.ptr(sx, sy), op.cfg.stride, w, h, cf, rf, mx, my, bd, pb)
        assert np.array_equal(d.all()[:h, :w], want), (k, w, h, cf, rf, mx, my)
        t = X.prep_8tap(I, rp, (sx, sy), w, h, cf, rf, mx, my, bd)
        wt = np.zeros((h, w), dtype=np.int16)
        L.r1o_prep_8tap(wt.ctypes.data, op.ptr(sx, sy), op.cfg.stride, w, h, cf, rf, mx, my, bd, pb)
        assert np.array_equal(t, wt), (k, "prep")
        t2 = np.clip(t.astype(np.int64) + rng.integers(-3000, 3000, size=t.shape), -32768, 32767).astype(np.int16)
        X.mc_avg(I, d, (0, 0), t, t2, w, h, bd)
        L.r1o_mc_avg(want.ctypes.data, w, t.ctypes.data, t2.ctypes.data, w, h, bd, pb)
        assert np.array_equal(d.all()[:h, :w], want), (k, "avg")
    assert I.calls["put_8tap_ref"] >= 24 and I.calls["get_filter"] >= 48 and I.calls["mc_avg_ref"] >= 24


def test_golden_files_hold_what_the_reference_text_computes(oracle):
    """spot re-computation of the committed golden vectors from the reference's text (a sample of every file)"""
    X = _X()
    g = load_golden("mc")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    for bd in (8, 12):
        I, rp, _ = _pair(X, oracle, bd, g["src_%d" % bd], xpad, ypad)
        off = 0
        for i, (sx, sy, w, h, cf, rf, mx, my) in enumerate(g["cases_%d" % bd]):
            w, h = int(w), int(h)
            if i % 9 == 0 and w * h <= 256:
                d = X.RPlane(I, w, h)
                X.put_8tap(I, d, (0, 0), rp, (int(sx), int(sy)), w, h, int(cf), int(rf), int(mx), int(my), bd)
                assert np.array_equal(d.all()[:h, :w].reshape(-1), g["put_%d" % bd][off:off + w * h]), i
                t = X.prep_8tap(I, rp, (int(sx), int(sy)), w, h, int(cf), int(rf), int(mx), int(my), bd)
                assert np.array_equal(t.reshape(-1), g["prep_%d" % bd][off:off + w * h]), i
            off += w * h
        po_x, po_y, w, h, mr, mc = (int(v) for v in g["pi_cases_%d" % bd][1])
        assert np.array_equal(X.predict_inter(I, rp, (po_x, po_y), w, h, (mr, mc), bd).reshape(-1),
                              g["pi_out_%d" % bd][64:64 + w * h])
    g = load_golden("me")
    I = X.interp("u8")
    for c, want in list(zip(g["range_cases"], g["range_out"]))[:20]:
        assert X.get_mv_range(I, *(int(v) for v in c)) == tuple(int(v) for v in want)
    for (ar, ac, br, bc, hp), want in list(zip(g["rate_cases"], g["rate_out"]))[::5]:
        assert X.get_mv_rate(I, (int(ar), int(ac)), (int(br), int(bc)), int(hp)) == int(want)
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    _, rorg, _ = _pair(X, oracle, 8, g["org_8"], xpad, ypad)
    _, rref, _ = _pair(X, oracle, 8, g["ref_8"], xpad, ypad)
    j = [int(v) for v in g["fs_jobs_8"][0]]
    m, c = X.full_search(I, rorg, rref, j[3], j[4], j[5], j[6], j[2], (j[0], j[1]), j[7], j[8], ((j[9], j[10]), (j[11], j[12])), j[13], 8)
    assert (m[0], m[1], c) == (int(g["fs_mv_8"][0][0]), int(g["fs_mv_8"][0][1]), int(g["fs_cost_8"][0]))
    k = 1
    j = [int(v) for v in g["sp_jobs_8"][k]]
    mr = tuple(int(v) for v in g["sp_range_8"][k])
    pm = ((j[4], j[5]), (j[6], j[7]))
    cand = tuple(int(v) for v in g["sp_cand_8"][k][0])
    assert X.get_mv_rd_cost(I, rorg, rref, (j[0], j[1]), j[2], cand, pm, j[3], j[8], j[9], mr, 8, True) == int(g["sp_cost_8"][k][0])
    got = X.telescopic_subpel_search(I, rorg, rref, (j[0], j[1]), j[2], j[3], pm, mr, j[8], j[9], 8, (j[10], j[11]),
                                     int(g["tele_start_cost_8"][k]))
    assert got == ((int(g["tele_mv_8"][k][0]), int(g["tele_mv_8"][k][1])), int(g["tele_cost_8"][k]))
    k = 0
    j = [int(v) for v in g["dia_jobs_8"][k]]
    preds = [tuple(int(v) for v in p) for p in g["dia_pred_8"][k]]
    got = X.diamond_me_search(I, rorg, rref, (j[0], j[1]), j[2], preds, ((j[4], j[5]), (j[6], j[7])), j[3], tuple(j[10:14]),
                              j[8], j[9], 8, bool(g["dia_sub_8"][k]))
    assert got == ((int(g["dia_mv_8"][k][0]), int(g["dia_mv_8"][k][1])), int(g["dia_cost_8"][k]))
    for name in ("full_search", "get_mv_rd_cost", "compute_mv_rd_cost", "telescopic_subpel_search", "diamond_me_search",
                 "get_best_predictor", "predict_inter", "get_params", "get_mv_rate", "get_mv_range"):
        assert I.calls.get(name, 0) > 0, name


def test_quant_and_rdo_golden_files_hold_what_the_reference_text_computes(oracle):
    X = _X()
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import make_ref_golden as M
    for name, gen in (("quant", M.gen_quant), ("rdo", M.gen_rdo)):
        g, fresh = load_golden(name), gen()
        assert sorted(g) == sorted(fresh), name
        for k in g:
            assert np.array_equal(g[k], fresh[k]), (name, k)
    I = X.interp("i32")
    for fn in ("update", "quantize", "dequantize", "divu_pair", "divu_gen", "get_log_tx_scale"):
        assert I.calls.get(fn, 0) > 0, fn
    I = X.interp("u16")
    for fn in ("sse_wxh", "cdef_dist_wxh", "cdef_dist_wxh_8x8", "compute_distortion_bias", "compute_mean_importance"):
        assert I.calls.get(fn, 0) > 0, fn


def test_quantize_text_vs_oracle_random(oracle):
    """QuantizationContext::quantize / dequantize (src/quantize.rs:257-334) on fresh inputs with the reference's own
    scan tables (src/scan_order.rs), incl. every TxType's scan of 8x8"""
    X = _X()
    I = X.interp("i32")
    rng = np.random.default_rng(31)
    for tt in range(16):
        ts = 1
        qi = int(rng.integers(0, 256))
        qc = X.quant_update(I, qi, ts, tt & 1, 8)
        scan = X.scan_order(I, ts, tt)
        assert sorted(scan.tolist()) == list(range(64))
        acq = int(qc.f["ac_quant"])
        co = rng.integers(-3 * acq, 3 * acq + 1, 64)
        q = X.quantize(I, qc, co, ts, tt)
        oq, orc = oracle.quantize_batch(oracle.quant_ctx(ts, tt & 1, int(qc.f["dc_quant"]), acq), ts, scan, co.reshape(1, 64),
                                        want_rcoeffs=True)
        assert np.array_equal(q, oq.reshape(-1)), tt
        assert np.array_equal(X.dequantize(I, qi, q, ts, 8), orc.reshape(-1)), tt
