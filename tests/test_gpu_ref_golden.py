"""The CUDA path, through the C ABI, against known answers made by executing the reference's own Rust text
(tests/golden/{mc,me,quant,rdo}_r02.npz; see tests/golden/make_ref_golden.py and tests/test_ref_golden.py, which
checks the CPU oracle against the same files).  Bit-exact.  Run on the B200 box: python -m pytest tests -m gpu"""
import numpy as np
import pytest

from golden_util import load_golden, oplane

pytestmark = pytest.mark.gpu
U64_MAX = (1 << 64) - 1


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


def dplane(r, oracle, ctx, visible, pb, xpad, ypad, xdec=0, ydec=0):
    o = oplane(oracle, visible, pb, xpad, ypad, xdec, ydec)
    h, w = visible.shape
    return o, r.Plane(ctx, w, h, xdec, ydec, xpad, ypad, pb).upload(o.data)


# ------------------------------------------------------------------------------------------------- MC ------
@pytest.mark.parametrize("bd", [8, 10, 12])
def test_put_prep_avg_match_reference(r, ctx, oracle, bd):
    g = load_golden("mc")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    o_src, d_src = dplane(r, oracle, ctx, g["src_%d" % bd], pb, xpad, ypad)
    cases = g["cases_%d" % bd]
    # prep: every case in ONE batch
    jobs = np.zeros(len(cases), dtype=r.MC_JOB)
    off = 0
    for i, (sx, sy, w, h, cf, rf, mx, my) in enumerate(cases):
        jobs[i] = (0, 0, sx, sy, w, h, cf, rf, mx, my, off)
        off += int(w) * int(h)
    tmp = np.full(off, -12345, dtype=np.int16)
    r.prep_8tap_batch(ctx, tmp, d_src, jobs, bd)
    assert np.array_equal(tmp, g["prep_%d" % bd])
    # put: each case into a fresh destination (blocks of one batch must not overlap), then a packed batch
    o_dst, d_dst = dplane(r, oracle, ctx, np.zeros((128, 512), dtype=np.uint16), pb, 8, 8)
    off = 0
    x = y = rowh = 0
    place = []
    for i, (sx, sy, w, h, cf, rf, mx, my) in enumerate(cases):
        w, h = int(w), int(h)
        if x + w > 512:
            x, y, rowh = 0, y + rowh, 0
        place.append((x, y))
        jobs[i]["dst_x"], jobs[i]["dst_y"] = x, y
        x += (w + 3) & ~3
        rowh = max(rowh, h)
    assert y + rowh <= 128
    r.put_8tap_batch(ctx, d_dst, d_src, jobs, bd)
    got = d_dst.download()
    c = o_dst.cfg
    for i, (sx, sy, w, h, cf, rf, mx, my) in enumerate(cases):
        w, h = int(w), int(h)
        px, py = place[i]
        blk = got[c.yorigin + py:c.yorigin + py + h, c.xorigin + px:c.xorigin + px + w]
        assert np.array_equal(blk.reshape(-1), g["put_%d" % bd][off:off + w * h]), ("put", i, tuple(cases[i]))
        off += w * h
    # mc_avg
    off = 0
    for w, h in g["avg_sizes_%d" % bd]:
        n = int(w) * int(h)
        o_d, d_d = dplane(r, oracle, ctx, np.zeros((32, 64), dtype=np.uint16), pb, 8, 8)
        r.mc_avg(ctx, d_d, (4, 8), np.ascontiguousarray(g["avg_t1_%d" % bd][off:off + n]),
                 np.ascontiguousarray(g["avg_t2_%d" % bd][off:off + n]), int(w), int(h), bd)
        out = d_d.download()
        blk = out[o_d.cfg.yorigin + 8:o_d.cfg.yorigin + 8 + int(h), o_d.cfg.xorigin + 4:o_d.cfg.xorigin + 4 + int(w)]
        assert np.array_equal(blk.reshape(-1), g["avg_out_%d" % bd][off:off + n]), ("avg", w, h)
        off += n


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_predict_inter_params_and_clamp_match_reference(r, ctx, oracle, bd):
    """r1c_mv_params (predict_inter's get_params + clamp, src/predict.rs:267-284) + put_8tap against whole blocks the
    reference's predict_inter produced, incl. MVs far outside the frame and a 4:2:0 chroma plane"""
    g = load_golden("mc")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    luma = dplane(r, oracle, ctx, g["src_%d" % bd], pb, xpad, ypad)
    chroma = dplane(r, oracle, ctx, g["csrc_%d" % bd], pb, 24, 24, 1, 1)
    for (o_p, d_p), ck, ok in ((luma, "pi_cases_%d", "pi_out_%d"), (chroma, "pc_cases_%d", "pc_out_%d")):
        off = 0
        for po_x, po_y, w, h, mr, mc in g[ck % bd]:
            w, h = int(w), int(h)
            rf, cf, sx, sy = r.mv_params(d_p.cfg, (int(po_x), int(po_y)), (int(mr), int(mc)))
            o_d, d_d = dplane(r, oracle, ctx, np.zeros((16, 16), dtype=np.uint16), pb, 8, 8)
            r.put_8tap(ctx, d_d, (0, 0), d_p, (sx, sy), w, h, cf, rf, 0, 0, bd)
            out = d_d.download()
            blk = out[o_d.cfg.yorigin:o_d.cfg.yorigin + h, o_d.cfg.xorigin:o_d.cfg.xorigin + w]
            assert np.array_equal(blk.reshape(-1), g[ok % bd][off:off + w * h]), (ck, po_x, po_y, w, h, mr, mc)
            off += w * h


# ------------------------------------------------------------------------------------------------- ME ------
def test_mv_range_and_rate_match_reference(r):
    g = load_golden("me")
    for c, want in zip(g["range_cases"], g["range_out"]):
        assert r.get_mv_range(int(c[0]), int(c[1]), (int(c[2]), int(c[3])), int(c[4]), int(c[5])) == tuple(int(v) for v in want)
    for (ar, ac, br, bc, hp), want in zip(g["rate_cases"], g["rate_out"]):
        assert r.get_mv_rate((int(ar), int(ac)), (int(br), int(bc)), int(hp)) == int(want)


def _fs_jobs(r, rows):
    jobs = np.zeros(len(rows), dtype=r.FS_JOB)
    for i, j in enumerate(rows):
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize"] = int(j[0]), int(j[1]), int(j[2])
        jobs[i]["x_lo"], jobs[i]["x_hi"], jobs[i]["y_lo"], jobs[i]["y_hi"] = (int(v) for v in j[3:7])
        jobs[i]["step"], jobs[i]["lambda_"] = int(j[7]), int(j[8])
        jobs[i]["pmv"][0] = (int(j[9]), int(j[10]))
        jobs[i]["pmv"][1] = (int(j[11]), int(j[12]))
        jobs[i]["allow_high_precision_mv"] = int(j[13])
    return jobs


@pytest.mark.parametrize("bd", [8, 10])
def test_full_search_matches_reference(r, ctx, oracle, bd):
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    _, d_org = dplane(r, oracle, ctx, g["org_%d" % bd], pb, xpad, ypad)
    _, d_ref = dplane(r, oracle, ctx, g["ref_%d" % bd], pb, xpad, ypad)
    res = r.full_search_batch(ctx, d_org, d_ref, _fs_jobs(r, g["fs_jobs_%d" % bd]), bd)
    assert np.array_equal(res["mv"]["row"], g["fs_mv_%d" % bd][:, 0])
    assert np.array_equal(res["mv"]["col"], g["fs_mv_%d" % bd][:, 1])
    assert np.array_equal(res["cost"], g["fs_cost_%d" % bd])


def test_full_search_ties_match_reference(r, ctx, oracle):
    g = load_golden("me")
    _, d = dplane(r, oracle, ctx, np.full((32, 32), 77, dtype=np.uint16), 1, 40, 40)
    res = r.full_search_batch(ctx, d, d, _fs_jobs(r, g["tie_jobs"]), 8)
    assert np.array_equal(res["mv"]["row"], g["tie_mv"][:, 0]) and np.array_equal(res["mv"]["col"], g["tie_mv"][:, 1])
    assert np.array_equal(res["cost"], g["tie_cost"])


def _sp_jobs(r, rows, ranges):
    jobs = np.zeros(len(rows), dtype=r.SUBPEL_JOB)
    for i, j in enumerate(rows):
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize"], jobs[i]["lambda_"] = (int(v) for v in j[:4])
        jobs[i]["pmv"][0] = (int(j[4]), int(j[5]))
        jobs[i]["pmv"][1] = (int(j[6]), int(j[7]))
        jobs[i]["use_satd"], jobs[i]["allow_high_precision_mv"] = int(j[8]), int(j[9])
        jobs[i]["mvx_min"], jobs[i]["mvx_max"], jobs[i]["mvy_min"], jobs[i]["mvy_max"] = (int(v) for v in ranges[i])
    return jobs


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_subpel_cost_and_telescopic_match_reference(r, ctx, oracle, bd):
    """get_mv_rd_cost through predict_inter (src/me.rs:787-856) per candidate, and r1c_subpel_refine_batch =
    motion_estimation's tail (:232-271): with use_satd it re-evaluates the start MV by SATD (the stored start cost of
    those jobs is that very value, made by the reference's get_mv_rd_cost), then telescopic_subpel_search (:858-941)"""
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    _, d_org = dplane(r, oracle, ctx, g["org_%d" % bd], pb, xpad, ypad)
    _, d_ref = dplane(r, oracle, ctx, g["ref_%d" % bd], pb, xpad, ypad)
    rows = g["sp_jobs_%d" % bd]
    jobs = _sp_jobs(r, rows, g["sp_range_%d" % bd])
    cand = np.zeros(g["sp_cand_%d" % bd].shape[:2], dtype=r.MV)
    cand["row"], cand["col"] = g["sp_cand_%d" % bd][:, :, 0], g["sp_cand_%d" % bd][:, :, 1]
    start = np.zeros(len(rows), dtype=r.ME_RESULT)
    start["mv"]["row"], start["mv"]["col"] = rows[:, 10], rows[:, 11]
    start["cost"] = g["tele_start_cost_%d" % bd]
    for hp in (0, 1):  # jobs of one batch share allow_high_precision_mv
        sel = np.nonzero(rows[:, 9] == hp)[0]
        if sel.size == 0:
            continue
        got = r.subpel_cost_batch(ctx, d_org, d_ref, jobs[sel], cand[sel], bd)
        assert np.array_equal(got, g["sp_cost_%d" % bd][sel]), hp
        out = r.subpel_refine_batch(ctx, d_org, d_ref, jobs[sel], start[sel], bd)
        assert np.array_equal(out["mv"]["row"], g["tele_mv_%d" % bd][sel, 0]), hp
        assert np.array_equal(out["mv"]["col"], g["tele_mv_%d" % bd][sel, 1]), hp
        assert np.array_equal(out["cost"], g["tele_cost_%d" % bd][sel]), hp


@pytest.mark.parametrize("bd", [8, 10])
def test_diamond_search_matches_reference(r, ctx, oracle, bd):
    g = load_golden("me")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    _, d_org = dplane(r, oracle, ctx, g["org_%d" % bd], pb, xpad, ypad)
    _, d_ref = dplane(r, oracle, ctx, g["ref_%d" % bd], pb, xpad, ypad)
    rows = g["dia_jobs_%d" % bd]
    jobs = _sp_jobs(r, rows, rows[:, 10:14])
    pred = np.zeros(g["dia_pred_%d" % bd].shape[:2], dtype=r.MV)
    pred["row"], pred["col"] = g["dia_pred_%d" % bd][:, :, 0], g["dia_pred_%d" % bd][:, :, 1]
    for i in range(len(rows)):
        out = r.diamond_search_batch(ctx, d_org, d_ref, jobs[i:i + 1], pred[i:i + 1], bd, subpixel=bool(g["dia_sub_%d" % bd][i]))
        assert (int(out["mv"]["row"][0]), int(out["mv"]["col"][0]), int(out["cost"][0])) == \
            (int(g["dia_mv_%d" % bd][i][0]), int(g["dia_mv_%d" % bd][i][1]), int(g["dia_cost_%d" % bd][i])), i


# --------------------------------------------------------------------------------------------- quantize ----
def test_quantize_dequantize_match_reference(r, ctx, oracle):
    g = load_golden("quant")
    for k, (ts, tt, qi, intra, bd, n) in enumerate(g["cases"]):
        c = [int(v) for v in g["ctx_%d" % k]]
        qc = r.QuantizationContext().update(ctx.lib, int(ts), int(intra), c[1], c[2])
        q = qc.c
        got = [q.log_tx_scale, q.dc_quant, q.ac_quant, q.dc_offset, q.ac_offset_eob, q.ac_offset0, q.ac_offset1] + \
            list(q.dc_mul_add) + list(q.ac_mul_add)
        assert got == c, (k, "r1c_quant_ctx_update")
        qq = qc.quantize(ctx, g["coeffs_%d" % k], g["scan_%d" % k])
        assert np.array_equal(qq, g["q_%d" % k]), (k, "quantize")
        assert np.array_equal(r.dequantize(ctx, g["q_%d" % k], int(ts), c[1], c[2]), g["r_%d" % k]), (k, "dequantize")


# -------------------------------------------------------------------------------------------------- RDO ----
@pytest.mark.parametrize("bd", [8, 10, 12])
def test_rdo_distortion_matches_reference(r, ctx, oracle, bd):
    g = load_golden("rdo")
    W, H, xpad, ypad = (int(v) for v in g["geom"])
    pb = 1 if bd == 8 else 2
    _, da = dplane(r, oracle, ctx, g["a_%d" % bd], pb, xpad, ypad)
    _, db = dplane(r, oracle, ctx, g["b_%d" % bd], pb, xpad, ypad)
    _, dca = dplane(r, oracle, ctx, g["a_%d" % bd][::2, ::2], pb, 8, 8, 1, 1)
    _, dcb = dplane(r, oracle, ctx, g["b_%d" % bd][::2, ::2], pb, 8, 8, 1, 1)
    imp = g["imp_%d" % bd]
    for (p1, p2), ck, ok in (((da, db), "cases_%d", "out_%d"), ((dca, dcb), "ccases_%d", "cout_%d")):
        for (x, y, w, h, kind, wb), want in zip(g[ck % bd], g[ok % bd]):
            job = np.zeros(1, dtype=r.RDO_JOB)
            job[0] = (x, y, x, y, w, h)
            kw = dict(block_importances=imp, w_in_imp_b=W // 8, w_in_b=W // 4, h_in_b=H // 4) if wb else {}
            got = r.sse_wxh(ctx, p1, p2, job, bd, cdef=bool(kind), **kw)
            assert int(got[0]) == int(want), (ck, x, y, w, h, kind, wb)
