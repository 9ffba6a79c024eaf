"""Writes tests/golden/{mc,me,quant,rdo}_r02.npz: known answers made by EXECUTING THE REFERENCE'S OWN RUST TEXT
(tests/ref_rs.py + tests/rustlite.py read /root/reference/src/*.rs at run time and interpret put_8tap_ref,
prep_8tap_ref, mc_avg_ref, predict_inter, get_mv_range, get_mv_rate, full_search, get_mv_rd_cost,
telescopic_subpel_search, diamond_me_search, QuantizationContext::update / quantize, dequantize, sse_wxh,
cdef_dist_wxh, compute_distortion_bias ...).  Dev container only (needs /root/reference); the .npz files travel to
the GPU box, where tests/test_ref_golden.py checks the oracle and tests/test_gpu_ref_golden.py the CUDA path
against them.  Nothing of the reference's source is stored: only inputs chosen here and the outputs it computed.

    python tests/golden/make_ref_golden.py [mc] [me] [quant] [rdo]

Array layouts are documented next to each writer; planes are stored as their visible pixels plus the
Plane::new arguments, and are rebuilt with Plane::new + pad on the consumer side.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_rs as X  # noqa: E402

PIXEL = {8: "u8", 10: "u16", 12: "u16"}
U64_MAX = (1 << 64) - 1


def smooth_field(rng, h, w, bd):
    v = rng.integers(0, 1 << bd, size=(h, w)).astype(np.float64)
    for _ in range(3):
        v = (v + np.roll(v, 1, 0) + np.roll(v, 1, 1) + np.roll(v, -1, 0) + np.roll(v, -1, 1)) / 5
    return np.clip(np.rint(v), 0, (1 << bd) - 1).astype(np.int64)


def rplane(I, vis, xdec=0, ydec=0, xpad=0, ypad=0, pad=True, pad_wh=None):
    h, w = vis.shape
    p = X.RPlane(I, w, h, xdec, ydec, xpad, ypad)
    p.set_visible(vis)
    if pad:
        if pad_wh:
            p.pad(*pad_wh)
        else:
            p.pad(w << xdec, h << ydec)
    return p


# ------------------------------------------------------------------------------------------------ MC --------
MC_SIZES = [(2, 2), (2, 4), (4, 2), (4, 4), (4, 8), (8, 4), (8, 8), (4, 16), (16, 4), (8, 16), (16, 8), (16, 16), (6, 10)]
MC_GEOM = dict(W=64, H=48, xpad=24, ypad=24)


def mc_cases(bd):
    """[n, 8] = src_x, src_y, w, h, col_frac, row_frac, mode_x, mode_y"""
    rng = np.random.default_rng(4000 + bd)
    W, H = MC_GEOM["W"], MC_GEOM["H"]
    cases, k = [], 0
    for mx in range(4):
        for my in range(4):
            for fc in range(4):
                w, h = MC_SIZES[k % len(MC_SIZES)]
                k += 1
                cf = 0 if fc in (0, 1) else int(rng.integers(1, 16))
                rf = 0 if fc in (0, 2) else int(rng.integers(1, 16))
                cases.append((int(rng.integers(-8, W - w + 9)), int(rng.integers(-8, H - h + 9)), w, h, cf, rf, mx, my))
    for f in range(1, 16):  # every fractional position on each axis
        cases.append((int(rng.integers(0, W - 8)), int(rng.integers(0, H - 8)), 8, 8, f, 0, 0, 0))
        cases.append((int(rng.integers(0, W - 8)), int(rng.integers(0, H - 8)), 8, 8, 0, f, 2, 2))
        cases.append((int(rng.integers(0, W - 4)), int(rng.integers(0, H - 4)), 4, 4, f, 16 - f, 1, 2))
    cases.append((5, 3, 32, 32, 7, 9, 0, 2))
    cases.append((-3, 20, 64, 16, 12, 5, 1, 0))
    cases.append((10, -5, 16, 32, 3, 14, 3, 3))
    return np.array(cases, dtype=np.int64)


def gen_mc():
    out = {}
    for bd in (8, 10, 12):
        I = X.interp(PIXEL[bd])
        rng = np.random.default_rng(4100 + bd)
        vis = rng.integers(0, 1 << bd, size=(MC_GEOM["H"], MC_GEOM["W"]))
        # extremes in one corner: the largest intermediates (i16 truncation of `as i16`, clamps)
        vis[:12, :12] = np.where(rng.random((12, 12)) < 0.5, 0, (1 << bd) - 1)
        src = rplane(I, vis, 0, 0, MC_GEOM["xpad"], MC_GEOM["ypad"])
        cases = mc_cases(bd)
        puts, preps = [], []
        for sx, sy, w, h, cf, rf, mx, my in cases:
            dst = X.RPlane(I, int(w), int(h))
            X.put_8tap(I, dst, (0, 0), src, (int(sx), int(sy)), int(w), int(h), int(cf), int(rf), int(mx), int(my), bd)
            puts.append(dst.all()[:h, :w].reshape(-1))
            preps.append(X.prep_8tap(I, src, (int(sx), int(sy)), int(w), int(h), int(cf), int(rf), int(mx), int(my), bd).reshape(-1))
        out["src_%d" % bd] = vis.astype(np.uint16)
        out["cases_%d" % bd] = cases
        out["put_%d" % bd] = np.concatenate(puts).astype(np.uint16)
        out["prep_%d" % bd] = np.concatenate(preps).astype(np.int16)
        # mc_avg: realistic prep outputs, then the extremes of i16
        avg_sizes = [(4, 4), (8, 8), (16, 8), (2, 2), (32, 4)]
        t1, t2, ao = [], [], []
        for i, (w, h) in enumerate(avg_sizes):
            if i < 3:
                a = preps[6 + i][:w * h] if preps[6 + i].size >= w * h else rng.integers(-500, 20000, w * h)
                a = np.resize(a, w * h).astype(np.int64)
                b = np.clip(a + rng.integers(-2000, 2000, w * h), -32768, 32767)
            else:
                a = rng.choice([-32768, 32767, 0, -1, 16383], w * h)
                b = rng.choice([-32768, 32767, 0, 1, 16384], w * h)
            dst = X.RPlane(I, w, h)
            X.mc_avg(I, dst, (0, 0), a, b, w, h, bd)
            t1.append(a)
            t2.append(b)
            ao.append(dst.all()[:h, :w].reshape(-1))
        out["avg_sizes_%d" % bd] = np.array(avg_sizes, dtype=np.int64)
        out["avg_t1_%d" % bd] = np.concatenate(t1).astype(np.int16)
        out["avg_t2_%d" % bd] = np.concatenate(t2).astype(np.int16)
        out["avg_out_%d" % bd] = np.concatenate(ao).astype(np.uint16)
        # predict_inter (MV -> offset / frac / clamp, src/predict.rs:267-284) incl. MVs far outside the frame and a
        # subsampled plane: [n, 6] = po_x, po_y, w, h, mv_row, mv_col
        pi_cases = [(8, 8, 8, 8, -13, 22), (0, 0, 8, 8, -300, -300), (56, 40, 8, 8, 500, 700), (16, 8, 4, 4, 7, -1),
                    (32, 16, 16, 8, -64, 33), (4, 40, 8, 8, 2000, -2000)]
        outs = []
        for po_x, po_y, w, h, mr, mc_ in pi_cases:
            outs.append(X.predict_inter(I, src, (po_x, po_y), w, h, (mr, mc_), bd).reshape(-1))
        out["pi_cases_%d" % bd] = np.array(pi_cases, dtype=np.int64)
        out["pi_out_%d" % bd] = np.concatenate(outs).astype(np.uint16)
        cvis = rng.integers(0, 1 << bd, size=(24, 32))
        csrc = rplane(I, cvis, 1, 1, 24, 24)
        pc_cases = [(4, 4, 4, 4, -13, 22), (0, 0, 8, 8, -301, 333), (24, 16, 8, 8, 1000, 1000), (8, 8, 2, 2, 5, 9)]
        outs = [X.predict_inter(I, csrc, (a, b), w, h, (mr, mc_), bd).reshape(-1) for a, b, w, h, mr, mc_ in pc_cases]
        out["csrc_%d" % bd] = cvis.astype(np.uint16)
        out["pc_cases_%d" % bd] = np.array(pc_cases, dtype=np.int64)
        out["pc_out_%d" % bd] = np.concatenate(outs).astype(np.uint16)
        print("mc", bd, "done", {k: v for k, v in I.calls.items() if k in ("put_8tap_ref", "prep_8tap_ref", "mc_avg_ref", "get_params")})
    out["geom"] = np.array([MC_GEOM["W"], MC_GEOM["H"], MC_GEOM["xpad"], MC_GEOM["ypad"]], dtype=np.int64)
    return out


# ------------------------------------------------------------------------------------------------ ME --------
ME_GEOM = dict(W=96, H=64, xpad=88, ypad=88)


def tdiv(a, b):
    q = abs(a) // b
    return q if a >= 0 else -q


def gen_me():
    out = {}
    I0 = X.interp("u8")
    rng = np.random.default_rng(5000)
    # get_mv_range: [n, 6] = w_in_b, h_in_b, bo_x, bo_y, blk_w, blk_h -> [n, 4]
    rc = []
    for _ in range(60):
        bw, bh = [(4, 4), (8, 8), (16, 16), (64, 64), (4, 16), (32, 8), (128, 128)][int(rng.integers(0, 7))]
        wb, hb = int(rng.integers(bw // 4, 500)), int(rng.integers(bh // 4, 300))
        rc.append((wb, hb, int(rng.integers(0, wb - bw // 4 + 1)), int(rng.integers(0, hb - bh // 4 + 1)), bw, bh))
    out["range_cases"] = np.array(rc, dtype=np.int64)
    out["range_out"] = np.array([X.get_mv_range(I0, *c) for c in rc], dtype=np.int64)
    # get_mv_rate: [n, 5] = a.row, a.col, b.row, b.col, allow_hp -> [n]
    mc_ = []
    for i in range(200):
        m = 4 if i < 50 else 2000 if i < 150 else 16000
        mc_.append((int(rng.integers(-m, m)), int(rng.integers(-m, m)), int(rng.integers(-m, m)), int(rng.integers(-m, m)), i & 1))
    mc_ += [(0, 0, 0, 0, 0), (1, 0, 0, 0, 0), (1, 0, 0, 0, 1), (-1, -1, 0, 0, 0), (16383, -16384, -16384, 16383, 1)]
    out["rate_cases"] = np.array(mc_, dtype=np.int64)
    out["rate_out"] = np.array([X.get_mv_rate(I0, (a, b), (c, d), e) for a, b, c, d, e in mc_], dtype=np.int64)

    W, H = ME_GEOM["W"], ME_GEOM["H"]
    for bd in (8, 10, 12):
        I = X.interp(PIXEL[bd])
        rng = np.random.default_rng(5100 + bd)
        rv = smooth_field(rng, H, W, bd)
        ov = np.clip(np.roll(rv, (2, -3), axis=(0, 1)) + rng.integers(-3, 4, size=(H, W)), 0, (1 << bd) - 1)
        ref = rplane(I, rv, 0, 0, ME_GEOM["xpad"], ME_GEOM["ypad"])
        org = rplane(I, ov, 0, 0, ME_GEOM["xpad"], ME_GEOM["ypad"])
        out["ref_%d" % bd] = rv.astype(np.uint16)
        out["org_%d" % bd] = ov.astype(np.uint16)
        t0 = time.time()
        if bd in (8, 10):
            # full_search: [n, 14] = po_x, po_y, bsize, x_lo, x_hi, y_lo, y_hi, step, lambda, pmv0.row, pmv0.col,
            # pmv1.row, pmv1.col, allow_hp -> [n, 3] = mv.row, mv.col, cost (as u64)
            jobs = []
            for k, (b, rx, ry, step) in enumerate([(0, 6, 5, 1), (3, 7, 6, 1), (6, 5, 5, 1), (16, 6, 4, 1), (17, 4, 6, 1),
                                                   (3, 8, 8, 2), (4, 5, 4, 1), (9, 3, 3, 1), (1, 9, 2, 1), (3, 6, 6, 1)]):
                bs = X.block_size(I, b)
                bw, bh = int(I.call("BlockSize::width", bs)), int(I.call("BlockSize::height", bs))
                if k in (1, 9):  # frame corners: the window is clipped by get_mv_range
                    po = (0, 0) if k == 1 else (W - bw, H - bh)
                else:
                    po = (int(rng.integers(0, (W - bw) // 4 + 1)) * 4, int(rng.integers(0, (H - bh) // 4 + 1)) * 4)
                mr = X.get_mv_range(I, W // 4, H // 4, po[0] // 4, po[1] // 4, bw, bh)
                cl = lambda v, lo, hi: min(max(v, lo), hi)  # noqa: E731
                x_lo = po[0] + cl(-rx, tdiv(mr[0], 8), tdiv(mr[1], 8))
                x_hi = po[0] + cl(rx, tdiv(mr[0], 8), tdiv(mr[1], 8))
                y_lo = po[1] + cl(-ry, tdiv(mr[2], 8), tdiv(mr[3], 8))
                y_hi = po[1] + cl(ry, tdiv(mr[2], 8), tdiv(mr[3], 8))
                jobs.append((po[0], po[1], b, x_lo, x_hi, y_lo, y_hi, step, [0, 40, 700, 5000][k % 4],
                             int(rng.integers(-40, 40)), int(rng.integers(-40, 40)), int(rng.integers(-40, 40)),
                             int(rng.integers(-40, 40)), k & 1))
            res = []
            for j in jobs:
                m, c = X.full_search(I, org, ref, j[3], j[4], j[5], j[6], j[2], (j[0], j[1]), j[7], j[8],
                                     ((j[9], j[10]), (j[11], j[12])), j[13], bd)
                res.append((m[0], m[1], c))
            out["fs_jobs_%d" % bd] = np.array(jobs, dtype=np.int64)
            out["fs_mv_%d" % bd] = np.array([r_[:2] for r_ in res], dtype=np.int64)
            out["fs_cost_%d" % bd] = np.array([r_[2] for r_ in res], dtype=np.uint64)
            print("me", bd, "full_search done %.1fs" % (time.time() - t0))
        # sub-pel jobs: [n, 12] = po_x, po_y, bsize, lambda, pmv0.row, pmv0.col, pmv1.row, pmv1.col, use_satd, allow_hp,
        # start.row, start.col ; mv range = get_mv_range of the block
        sj = []
        for k, b in enumerate([3, 0, 6, 16, 1, 3] if bd != 12 else [3, 0, 4]):
            bs = X.block_size(I, b)
            bw, bh = int(I.call("BlockSize::width", bs)), int(I.call("BlockSize::height", bs))
            po = (int(rng.integers(0, (W - bw) // 4 + 1)) * 4, int(rng.integers(0, (H - bh) // 4 + 1)) * 4)
            if k == 5:
                po = (0, 0)
            sj.append((po[0], po[1], b, [0, 40, 900][k % 3], int(rng.integers(-30, 30)), int(rng.integers(-30, 30)),
                       int(rng.integers(-30, 30)), int(rng.integers(-30, 30)), 1 if k % 3 else 0, k & 1,
                       int(rng.integers(-2, 3)) * 8, int(rng.integers(-2, 3)) * 8))
        ranges, costs, tele, cands = [], [], [], []
        for k, j in enumerate(sj):
            bs = X.block_size(I, j[2])
            bw, bh = int(I.call("BlockSize::width", bs)), int(I.call("BlockSize::height", bs))
            mr = X.get_mv_range(I, W // 4, H // 4, j[0] // 4, j[1] // 4, bw, bh)
            ranges.append(mr)
            pm = ((j[4], j[5]), (j[6], j[7]))
            cc = [(j[10] + int(rng.integers(-9, 10)), j[11] + int(rng.integers(-9, 10))) for _ in range(3)]
            if k == 5:
                cc[0] = (mr[2] - 1, mr[0])  # one step outside the legal range -> u64::MAX
                cc[1] = (mr[2], mr[0])      # the corner of the legal range: the clamp of predict_inter is exercised
            cands.append(cc)
            costs.append([X.get_mv_rd_cost(I, org, ref, (j[0], j[1]), j[2], c, pm, j[3], j[8], j[9], mr, bd, True) for c in cc])
            start_cost = [U64_MAX, 30000, 0][k % 3]
            if j[8]:  # use_satd: the start cost motion_estimation hands over, get_mv_rd_cost full-pel (src/me.rs:235-252)
                start_cost = X.get_mv_rd_cost(I, org, ref, (j[0], j[1]), j[2], (j[10], j[11]), pm, j[3], 1, j[9], mr, bd, False)
            tele.append(X.telescopic_subpel_search(I, org, ref, (j[0], j[1]), j[2], j[3], pm, mr, j[8], j[9], bd,
                                                   (j[10], j[11]), start_cost) + (start_cost,))
        out["sp_jobs_%d" % bd] = np.array(sj, dtype=np.int64)
        out["sp_range_%d" % bd] = np.array(ranges, dtype=np.int64)
        out["sp_cand_%d" % bd] = np.array(cands, dtype=np.int64)
        out["sp_cost_%d" % bd] = np.array(costs, dtype=np.uint64)
        out["tele_mv_%d" % bd] = np.array([m for m, c, s_ in tele], dtype=np.int64)
        out["tele_cost_%d" % bd] = np.array([c for m, c, s_ in tele], dtype=np.uint64)
        out["tele_start_cost_%d" % bd] = np.array([s_ for m, c, s_ in tele], dtype=np.uint64)
        print("me", bd, "sub-pel done %.1fs" % (time.time() - t0))
        if bd in (8, 10):
            # diamond_me_search: jobs as above + 5 predictors [n, 5, 2] + subpixel flag [n] -> [n, 3]
            dj, dp, dsub, dres = [], [], [], []
            for k, b in enumerate([3, 6, 0, 4]):
                bs = X.block_size(I, b)
                bw, bh = int(I.call("BlockSize::width", bs)), int(I.call("BlockSize::height", bs))
                po = (int(rng.integers(0, (W - bw) // 4 + 1)) * 4, int(rng.integers(0, (H - bh) // 4 + 1)) * 4)
                mr = X.get_mv_range(I, W // 4, H // 4, po[0] // 4, po[1] // 4, bw, bh)
                j = (po[0], po[1], b, [0, 40, 900][k % 3], int(rng.integers(-30, 30)), int(rng.integers(-30, 30)),
                     int(rng.integers(-30, 30)), int(rng.integers(-30, 30)), k & 1, (k >> 1) & 1) + mr
                preds = [(0, 0)] + [(int(rng.integers(-60, 61)), int(rng.integers(-60, 61))) for _ in range(3)]
                preds.append((mr[3] + 8, 0) if k == 1 else preds[1])
                sub = k >= 2
                m, c = X.diamond_me_search(I, org, ref, (po[0], po[1]), b, preds, ((j[4], j[5]), (j[6], j[7])), j[3], mr,
                                           j[8], j[9], bd, sub)
                dj.append(j)
                dp.append(preds)
                dsub.append(int(sub))
                dres.append((m[0], m[1], c))
            out["dia_jobs_%d" % bd] = np.array(dj, dtype=np.int64)
            out["dia_pred_%d" % bd] = np.array(dp, dtype=np.int64)
            out["dia_sub_%d" % bd] = np.array(dsub, dtype=np.int64)
            out["dia_mv_%d" % bd] = np.array([r_[:2] for r_ in dres], dtype=np.int64)
            out["dia_cost_%d" % bd] = np.array([r_[2] for r_ in dres], dtype=np.uint64)
            print("me", bd, "diamond done %.1fs" % (time.time() - t0))
    # ties: a flat plane pair, every candidate has SAD 0 -> rate and raster order decide (strict `<`)
    I = X.interp("u8")
    flat = np.full((32, 32), 77, dtype=np.int64)
    fp = rplane(I, flat, 0, 0, 40, 40)
    tj = [(12, 12, 3, 8, 16, 9, 15, 1, 0, 0, 0, 0, 0, 0), (12, 12, 3, 8, 16, 9, 15, 1, 100, 5, -9, 30, 30, 0),
          (8, 8, 0, 4, 12, 4, 12, 2, 100, 0, 0, 0, 0, 1)]
    res = []
    for j in tj:
        m, c = X.full_search(I, fp, fp, j[3], j[4], j[5], j[6], j[2], (j[0], j[1]), j[7], j[8], ((j[9], j[10]), (j[11], j[12])), j[13], 8)
        res.append((m[0], m[1], c))
    out["tie_jobs"] = np.array(tj, dtype=np.int64)
    out["tie_mv"] = np.array([r_[:2] for r_ in res], dtype=np.int64)
    out["tie_cost"] = np.array([r_[2] for r_ in res], dtype=np.uint64)
    out["geom"] = np.array([W, H, ME_GEOM["xpad"], ME_GEOM["ypad"]], dtype=np.int64)
    print({k: v for k, v in I.calls.items() if k in ("full_search", "get_mv_rd_cost", "telescopic_subpel_search", "diamond_me_search", "predict_inter")})
    return out


# ---------------------------------------------------------------------------------------------- quantize ----
QUANT_CASES = [  # tx_size, tx_type, qindex, is_intra, bit_depth, n_blocks
    (0, 0, 40, 0, 8, 6), (0, 9, 200, 1, 8, 4), (1, 0, 100, 0, 8, 6), (1, 1, 100, 1, 10, 4), (1, 10, 30, 0, 12, 3),
    (2, 0, 128, 0, 8, 3), (5, 0, 60, 1, 8, 4), (14, 11, 90, 0, 10, 3), (3, 0, 110, 0, 8, 2), (9, 0, 150, 1, 10, 1),
    (4, 0, 100, 0, 8, 1), (17, 0, 80, 0, 10, 1), (11, 0, 255, 0, 8, 1), (1, 0, 0, 0, 8, 2), (1, 0, 255, 1, 12, 2)]


def gen_quant():
    I = X.interp("i32")
    out = {"cases": np.array(QUANT_CASES, dtype=np.int64)}
    rng = np.random.default_rng(6000)
    for k, (ts, tt, qi, intra, bd, n) in enumerate(QUANT_CASES):
        t0 = time.time()
        tse = X.tx_size_enum(I, ts)
        w, h = int(I.call("TxSize::width", tse)), int(I.call("TxSize::height", tse))
        area = w * h
        qc = X.quant_update(I, qi, ts, intra, bd)
        f = qc.f
        out["ctx_%d" % k] = np.array([int(f["log_tx_scale"]), int(f["dc_quant"]), int(f["ac_quant"]), int(f["dc_offset"]),
                                      int(f["ac_offset_eob"]), int(f["ac_offset0"]), int(f["ac_offset1"])] +
                                     [int(x) for x in f["dc_mul_add"]] + [int(x) for x in f["ac_mul_add"]], dtype=np.int64)
        scan = X.scan_order(I, ts, tt)
        out["scan_%d" % k] = scan
        coded = scan.size
        co = np.zeros((n, area), dtype=np.int64)
        acq = int(f["ac_quant"])
        for b in range(n):
            kind = (b + k) % 4
            mag = 1 << (bd + 3)
            if kind == 0:  # decaying, realistic
                env = mag / (1.0 + np.arange(area) * 16.0 / area) ** 3
                c = (rng.standard_normal(area) * env).astype(np.int64)
                c[scan.astype(np.int64)] = np.sort(np.abs(c[scan.astype(np.int64)]))[::-1] * rng.choice([-1, 1], coded)
            elif kind == 1:  # around the rounding thresholds: |c| in 0 .. 3 quantizer steps
                c = rng.integers(-3 * acq, 3 * acq + 1, area)
            elif kind == 2:  # dense and large
                c = rng.integers(-mag, mag + 1, area)
            else:  # sparse
                c = np.where(rng.random(area) < 0.1, rng.integers(-4 * acq, 4 * acq + 1, area), 0)
            co[b] = c
        q = np.stack([X.quantize(I, qc, co[b], ts, tt) for b in range(n)])
        r = np.stack([X.dequantize(I, qi, q[b][:coded], ts, bd) for b in range(n)])
        assert not q[:, coded:].any()
        out["coeffs_%d" % k] = co.astype(np.int32)
        out["q_%d" % k] = q[:, :coded].astype(np.int32)
        out["r_%d" % k] = r.astype(np.int32)
        print("quant case", k, (ts, tt, qi, intra, bd, n), "%.1fs" % (time.time() - t0))
    # divu_pair: [n, 2] = x, d -> x / d
    dv = [(int(rng.integers(-(1 << 26), 1 << 26)), int(rng.choice([1, 2, 3, 4, 7, 8, 93, 112, 255, 1000, 1336, 21387])))
          for _ in range(300)]
    out["divu_cases"] = np.array(dv, dtype=np.int64)
    out["divu_out"] = np.array([X.divu_pair(I, x, d) for x, d in dv], dtype=np.int64)
    print({k: v for k, v in I.calls.items() if k in ("update", "quantize", "dequantize", "divu_pair", "divu_gen")})
    return out


# ------------------------------------------------------------------------------------------------- RDO ------
RDO_GEOM = dict(W=64, H=48, xpad=16, ypad=16)


def gen_rdo():
    out = {}
    W, H = RDO_GEOM["W"], RDO_GEOM["H"]
    for bd in (8, 10, 12):
        I = X.interp(PIXEL[bd])
        rng = np.random.default_rng(7000 + bd)
        av = smooth_field(rng, H, W, bd)
        bv = np.clip(av + rng.integers(-(1 << (bd - 4)), 1 << (bd - 4), size=(H, W)), 0, (1 << bd) - 1)
        bv[:8, :8] = (1 << bd) - 1 - av[:8, :8]  # one block of extreme differences
        a = rplane(I, av, 0, 0, RDO_GEOM["xpad"], RDO_GEOM["ypad"])
        b = rplane(I, bv, 0, 0, RDO_GEOM["xpad"], RDO_GEOM["ypad"])
        imp = (rng.random((H // 8) * (W // 8)) * 6).astype(np.float32)
        imp[:3] = [0.0, 1e-3, 37.5]
        fi = X.frame_invariants(I, bd, False, w_in_b=W // 4, h_in_b=H // 4, block_importances=imp, w_in_imp_b=W // 8)
        # [n, 6] = x, y, w, h, kind (0 sse, 1 cdef), with_bias -> u64
        cases = [(0, 0, 8, 8, 0, 1), (0, 0, 8, 8, 1, 1), (8, 16, 16, 16, 0, 1), (8, 16, 16, 16, 1, 1), (16, 8, 32, 16, 0, 0),
                 (16, 8, 32, 16, 1, 0), (4, 4, 4, 4, 0, 1), (12, 20, 4, 8, 0, 1), (32, 16, 32, 32, 0, 1), (32, 16, 32, 32, 1, 1),
                 (56, 40, 8, 8, 0, 1), (56, 40, 8, 8, 1, 1), (0, 24, 64, 16, 0, 1), (40, 0, 16, 4, 0, 0)]
        res = []
        for x, y, w, h, kind, wb in cases:
            if kind == 0:
                res.append(X.sse_wxh(I, a, (x, y), b, (x, y), w, h, fi if wb else None))
            else:
                res.append(X.cdef_dist_wxh(I, a, (x, y), b, (x, y), w, h, bd, fi if wb else None))
        out["a_%d" % bd] = av.astype(np.uint16)
        out["b_%d" % bd] = bv.astype(np.uint16)
        out["imp_%d" % bd] = imp
        out["cases_%d" % bd] = np.array(cases, dtype=np.int64)
        out["out_%d" % bd] = np.array(res, dtype=np.uint64)
        # chroma plane (xdec = ydec = 1): importance blocks are 4x4 pixels there
        cav, cbv = av[::2, ::2].copy(), bv[::2, ::2].copy()
        ca = rplane(I, cav, 1, 1, 8, 8)
        cb = rplane(I, cbv, 1, 1, 8, 8)
        ccases = [(0, 0, 4, 4, 0, 1), (4, 8, 8, 8, 0, 1), (16, 8, 16, 8, 0, 1), (8, 4, 4, 8, 0, 0)]
        out["ccases_%d" % bd] = np.array(ccases, dtype=np.int64)
        out["cout_%d" % bd] = np.array([X.sse_wxh(I, ca, (x, y), cb, (x, y), w, h, fi if wb else None)
                                        for x, y, w, h, _k, wb in ccases], dtype=np.uint64)
        print("rdo", bd, {k: v for k, v in I.calls.items() if k in ("sse_wxh", "cdef_dist_wxh_8x8", "compute_distortion_bias")})
    out["geom"] = np.array([W, H, RDO_GEOM["xpad"], RDO_GEOM["ypad"]], dtype=np.int64)
    return out


GENERATORS = {"mc": gen_mc, "me": gen_me, "quant": gen_quant, "rdo": gen_rdo}

if __name__ == "__main__":
    assert X.available(), "needs /root/reference"
    for name in (sys.argv[1:] or list(GENERATORS)):
        t = time.time()
        d = GENERATORS[name]()
        path = os.path.join(HERE, "%s_r02.npz" % name)
        np.savez_compressed(path, **d)
        print("wrote", path, os.path.getsize(path), "bytes in %.1fs" % (time.time() - t))
