"""Seeded cases behind tests/golden/oracle_r01.npz.

`oracle_vectors(r1o)` computes every vector with the CPU oracle (this is what make_fixtures.py
stored); `cuda_vectors(r, ctx, r1o)` computes the same vectors through the C ABI.  Inputs are
rebuilt from fixed seeds, only outputs are stored (full arrays when small, else the first 8 bytes
of a SHA-256 as a u64)."""
import hashlib
import os
import sys

import numpy as np

BDS = [(1, 8), (2, 10), (2, 12)]
FILTER_PAIRS = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (1, 2), (2, 0), (2, 1), (2, 2), (3, 3)]
MC_SIZES = [(4, 4), (8, 4), (4, 16), (8, 8), (16, 8), (32, 16), (64, 64), (128, 64)]
FRAME = dict(width=128, height=96, range_x=8, range_y=8, lambda_=160)


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest()[:8], dtype="<u8")[0]


def plane(r1o, w, h, pb, bd, seed, xpad=88, ypad=88, smooth=False):
    rng = np.random.default_rng(seed)
    o = r1o.OPlane(w, h, 0, 0, xpad, ypad, pb)
    v = rng.integers(0, 1 << bd, size=(h, w))
    if smooth:
        v = v.astype(np.float64)
        for _ in range(3):
            v = (v + np.roll(v, 1, 0) + np.roll(v, 1, 1) + np.roll(v, -1, 0) + np.roll(v, -1, 1)) / 5
        v = np.clip(np.rint(v), 0, (1 << bd) - 1)
    o.visible[...] = v.astype(o.dtype)
    o.pad()
    return o


def dist_jobs(r1o, seed):
    """one job per BlockSize (all 22) at seeded positions of a 320 x 200 plane pair"""
    rng = np.random.default_rng(seed)
    L = r1o.lib()
    out = []
    for b in range(22):
        w, h = L.r1o_block_width(b), L.r1o_block_height(b)
        out.append((int(rng.integers(0, 320 - w)), int(rng.integers(0, 200 - h)), int(rng.integers(-40, 320 - w + 40)),
                    int(rng.integers(-40, 200 - h + 40)), b))
    return out


def mc_jobs(seed):
    """(w, h) blocks read from a 256 x 128 source plane and written to disjoint places of a 1024 x 128
    destination (one 128-px band per size), seeded fracs / filter pairs / offsets; every 4th job has a
    zero col_frac, every 3rd a zero row_frac -> all four cases"""
    rng = np.random.default_rng(seed)
    jobs = []
    for band, (w, h) in enumerate(MC_SIZES):
        for y in range(0, 128 - h + 1, max(h, 32)):
            for x in range(0, 128 - w + 1, max(w, 64)):
                k = len(jobs)
                mx, my = FILTER_PAIRS[int(rng.integers(0, 10))]
                jobs.append((band * 128 + x, y, x + int(rng.integers(-20, 21)) + 64, y + int(rng.integers(-20, 21)), w, h,
                             0 if k % 4 == 0 else int(rng.integers(1, 16)), 0 if k % 3 == 0 else int(rng.integers(1, 16)),
                             mx, my))
    return jobs


def tx_cases(r1o):
    L = r1o.lib()
    return [(ts, tt) for ts in range(19) for tt in range(16) if L.r1o_fht_valid(ts, tt)]


def residual(r1o, ts, tt, bd):
    L = r1o.lib()
    rng = np.random.default_rng(7000 + ts * 16 + tt + 1000 * bd)
    lim = (1 << bd) - 1
    return rng.integers(-lim, lim + 1, size=(4, L.r1o_tx_height(ts) * L.r1o_tx_width(ts))).astype(np.int16)


def rdo_jobs(seed, mult):
    rng = np.random.default_rng(seed)
    jobs = []
    for _ in range(40):
        w, h = mult << int(rng.integers(0, 4)), mult << int(rng.integers(0, 4))
        jobs.append((int(rng.integers(0, (192 - w) // mult + 1)) * mult, int(rng.integers(0, (128 - h) // mult + 1)) * mult, w, h))
    return jobs


def importances(seed=4242):
    return (np.random.default_rng(seed).random(24 * 16) * 4).astype(np.float32)  # 192x128 luma: 48x32 MI -> 24x16


# ------------------------------------------------------------------------- oracle side
def oracle_vectors(r1o):
    import ctypes as C
    L = r1o.lib()
    out = {}
    for pb, bd in BDS:
        org, ref = plane(r1o, 320, 200, pb, bd, 11 + bd), plane(r1o, 320, 200, pb, bd, 12 + bd, 136, 100)
        jobs = dist_jobs(r1o, 13)
        out["sad_%d" % bd] = np.array([r1o.sad(org, j[0], j[1], ref, j[2], j[3], L.r1o_block_width(j[4]), L.r1o_block_height(j[4]))
                                       for j in jobs], dtype=np.uint32)
        out["satd_%d" % bd] = np.array([r1o.satd(org, j[0], j[1], ref, j[2], j[3], L.r1o_block_width(j[4]), L.r1o_block_height(j[4]))
                                        for j in jobs], dtype=np.uint32)
        # MC
        src = plane(r1o, 256, 128, pb, bd, 21 + bd)
        dst = r1o.OPlane(1024, 128, 0, 0, 88, 88, pb)
        jobs = mc_jobs(22)
        total = sum(j[4] * j[5] for j in jobs)
        t1, t2 = np.zeros(total, np.int16), np.zeros(total, np.int16)
        off = 0
        for j in jobs:
            x, y, sx, sy, w, h, cf, rf, mx, my = j
            L.r1o_put_8tap(dst.ptr(x, y), dst.cfg.stride, src.ptr(sx, sy), src.cfg.stride, w, h, cf, rf, mx, my, bd, pb)
            L.r1o_prep_8tap(t1[off:].ctypes.data, src.ptr(sx, sy), src.cfg.stride, w, h, cf, rf, mx, my, bd, pb)
            L.r1o_prep_8tap(t2[off:].ctypes.data, src.ptr(sx + 1, sy + 2), src.cfg.stride, w, h, rf, cf, my, mx, bd, pb)
            off += w * h
        out["mc_put_%d" % bd] = digest(dst.data)
        out["mc_prep_%d" % bd] = digest(t1)
        avg = r1o.OPlane(1024, 128, 0, 0, 88, 88, pb)
        off = 0
        for j in jobs:
            x, y, _, _, w, h = j[:6]
            L.r1o_mc_avg(avg.ptr(x, y), avg.cfg.stride, t1[off:].ctypes.data, t2[off:].ctypes.data, w, h, bd, pb)
            off += w * h
        out["mc_avg_%d" % bd] = digest(avg.data)
        # forward transforms: one digest per valid (size, type)
        out["fht_%d" % bd] = np.array([digest(r1o.fht_batch(residual(r1o, ts, tt, bd), ts, tt, bd)) for ts, tt in tx_cases(r1o)],
                                      dtype=np.uint64)
        # RDO distortion
        a, b = plane(r1o, 192, 128, pb, bd, 31 + bd, 16, 16), plane(r1o, 192, 128, pb, bd, 32 + bd, 48, 32)
        imp = importances()
        out["sse_%d" % bd] = np.array([r1o.sse_wxh(a, x, y, b, x, y, w, h, imp, 24, 48, 32) for x, y, w, h in rdo_jobs(33, 4)], dtype=np.uint64)
        out["cdef_%d" % bd] = np.array([r1o.cdef_dist_wxh(a, x, y, b, x, y, w, h, bd, imp, 24, 48, 32) for x, y, w, h in rdo_jobs(34, 8)],
                                       dtype=np.uint64)
    # whole-frame motion estimation (8-bit and 10-bit)
    for pb, bd in BDS[:2]:
        ref = plane(r1o, FRAME["width"], FRAME["height"], pb, bd, 41 + bd, smooth=True)
        org = frame_org(r1o, ref, pb, bd)
        mv, cost = r1o.frame_me(org, ref, FRAME["range_x"], FRAME["range_y"], FRAME["lambda_"], ALL_SIZES_TO_64, 1, 1, 0, bd)
        out["frame_mv_%d" % bd] = np.asarray(mv).reshape(-1, 2).astype(np.int16)
        out["frame_cost_%d" % bd] = np.asarray(cost, dtype=np.uint64)
    rng = np.random.default_rng(51)
    out["tx_dist"] = np.array([r1o.tx_dist(rng.integers(-(1 << 17), 1 << 17, L.r1o_tx_width(ts) * L.r1o_tx_height(ts)),
                                           rng.integers(-(1 << 17), 1 << 17, min(L.r1o_tx_width(ts), 32) * min(L.r1o_tx_height(ts), 32)), ts)
                               for ts in range(19)], dtype=np.uint64)
    for name, (q, rc) in quant_cases(r1o, lambda qc, ts, scan, co, isi, dcq, acq: r1o.quantize_batch(
            r1o.quant_ctx(ts, isi, dcq, acq), ts, scan, co, want_rcoeffs=True)).items():
        out["quant_q_" + name], out["quant_r_" + name] = digest(q), digest(rc)
    del C
    return out


def quant_cases(r1o, run):
    """quantize + dequantize of seeded blocks of every TxSize (zig-zag scan, inter and intra)"""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from quant_util import synth_coeffs, zigzag_scan
    L = r1o.lib()
    rng = np.random.default_rng(61)
    res = {}
    for ts in range(19):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        scan = zigzag_scan(min(w, 32), min(h, 32))
        for isi, (dcq, acq) in ((0, (37, 41)), (1, (1336, 1828))):
            co = synth_coeffs(rng, 9, w * h, "decay", acq)
            res["%d_%d" % (ts, isi)] = run(None, ts, scan, co, isi, dcq, acq)
    return res


ALL_SIZES_TO_64 = sum(1 << b for b in range(22) if b not in (13, 14, 15))


def frame_org(r1o, ref, pb, bd):
    rng = np.random.default_rng(43)
    v = np.roll(ref.visible, (2, -3), axis=(0, 1)).astype(np.int64) + rng.integers(-2, 3, size=ref.visible.shape)
    o = r1o.OPlane(FRAME["width"], FRAME["height"], 0, 0, 88, 88, pb)
    o.visible[...] = np.clip(v, 0, (1 << bd) - 1).astype(o.dtype)
    o.pad()
    return o


# --------------------------------------------------------------------------- CUDA side
def cuda_vectors(r, ctx, r1o):
    """the same vectors through the C ABI (rav1e_b200.api); planes are uploaded from the seeded host planes"""
    L = r1o.lib()
    out = {}

    def dev(o):
        c = o.cfg
        return r.Plane(ctx, int(c.width), int(c.height), int(c.xdec), int(c.ydec), int(c.xpad), int(c.ypad), o.px_bytes).upload(o.data)

    for pb, bd in BDS:
        org, ref = plane(r1o, 320, 200, pb, bd, 11 + bd), plane(r1o, 320, 200, pb, bd, 12 + bd, 136, 100)
        jobs = np.array(dist_jobs(r1o, 13), dtype=r.DIST_JOB)
        d_org, d_ref = dev(org), dev(ref)
        out["sad_%d" % bd] = r.sad_batch(ctx, d_org, d_ref, jobs, bd)
        out["satd_%d" % bd] = r.satd_batch(ctx, d_org, d_ref, jobs, bd)
        src = plane(r1o, 256, 128, pb, bd, 21 + bd)
        d_src = dev(src)
        d_dst = dev(r1o.OPlane(1024, 128, 0, 0, 88, 88, pb))
        cj = mc_jobs(22)
        total = sum(j[4] * j[5] for j in cj)
        mj, mj2, aj = np.zeros(len(cj), dtype=r.MC_JOB), np.zeros(len(cj), dtype=r.MC_JOB), np.zeros(len(cj), dtype=r.AVG_JOB)
        off = 0
        for i, (x, y, sx, sy, w, h, cf, rf, mx, my) in enumerate(cj):
            mj[i] = (x, y, sx, sy, w, h, cf, rf, mx, my, off)
            mj2[i] = (x, y, sx + 1, sy + 2, w, h, rf, cf, my, mx, off)
            aj[i] = (x, y, w, h, off)
            off += w * h
        r.put_8tap_batch(ctx, d_dst, d_src, mj, bd)
        out["mc_put_%d" % bd] = digest(d_dst.download())
        t1, t2 = np.zeros(total, np.int16), np.zeros(total, np.int16)
        r.prep_8tap_batch(ctx, t1, d_src, mj, bd)
        r.prep_8tap_batch(ctx, t2, d_src, mj2, bd)
        out["mc_prep_%d" % bd] = digest(t1)
        d_avg = dev(r1o.OPlane(1024, 128, 0, 0, 88, 88, pb))
        r.mc_avg_batch(ctx, d_avg, t1, t2, aj, bd)
        out["mc_avg_%d" % bd] = digest(d_avg.download())
        out["fht_%d" % bd] = np.array([digest(r.forward_transform(ctx, residual(r1o, ts, tt, bd), ts, tt, bd)) for ts, tt in tx_cases(r1o)],
                                      dtype=np.uint64)
        a, b = plane(r1o, 192, 128, pb, bd, 31 + bd, 16, 16), plane(r1o, 192, 128, pb, bd, 32 + bd, 48, 32)
        d_a, d_b = dev(a), dev(b)
        imp = importances()
        j = np.array([(x, y, x, y, w, h) for x, y, w, h in rdo_jobs(33, 4)], dtype=r.RDO_JOB)
        out["sse_%d" % bd] = r.sse_wxh(ctx, d_a, d_b, j, bd, imp, 24, 48, 32)
        j = np.array([(x, y, x, y, w, h) for x, y, w, h in rdo_jobs(34, 8)], dtype=r.RDO_JOB)
        out["cdef_%d" % bd] = r.cdef_dist_wxh(ctx, d_a, d_b, j, bd, imp, 24, 48, 32)
    for pb, bd in BDS[:2]:
        ref = plane(r1o, FRAME["width"], FRAME["height"], pb, bd, 41 + bd, smooth=True)
        org = frame_org(r1o, ref, pb, bd)
        p = r.make_frame_me_params(FRAME["width"], FRAME["height"], FRAME["range_x"], FRAME["range_y"], FRAME["lambda_"])
        res = r.frame_me(ctx, dev(org), dev(ref), p, bd)
        out["frame_mv_%d" % bd] = np.stack([res["mv"]["row"], res["mv"]["col"]], axis=1).astype(np.int16)
        out["frame_cost_%d" % bd] = res["cost"].astype(np.uint64)
    rng = np.random.default_rng(51)
    td = []
    for ts in range(19):
        c = rng.integers(-(1 << 17), 1 << 17, L.r1o_tx_width(ts) * L.r1o_tx_height(ts))
        rc = rng.integers(-(1 << 17), 1 << 17, min(L.r1o_tx_width(ts), 32) * min(L.r1o_tx_height(ts), 32))
        td.append(r.tx_distortion(ctx, c, rc, ts)[0])
    out["tx_dist"] = np.array(td, dtype=np.uint64)

    def run(_, ts, scan, co, isi, dcq, acq):
        q = r.QuantizationContext().update(ctx.lib, ts, isi, dcq, acq).quantize(ctx, co, scan)
        return q, r.dequantize(ctx, q, ts, dcq, acq)
    for name, (q, rc) in quant_cases(r1o, run).items():
        out["quant_q_" + name], out["quant_r_" + name] = digest(q), digest(rc)
    return out
