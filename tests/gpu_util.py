"""Shared helpers of the GPU parity tests: twin planes (oracle host plane +
device plane with identical bytes) and oracle-side reference loops."""
import ctypes as C

import numpy as np

U64_MAX = (1 << 64) - 1


def twin_plane(r, r1o, ctx, width, height, px_bytes, bit_depth, rng=None, xpad=88, ypad=88,
               xdec=0, ydec=0, data=None, pad=True, smooth=False):
    """-> (oracle OPlane, device Plane) holding the same padded content"""
    o = r1o.OPlane(width, height, xdec, ydec, xpad, ypad, px_bytes)
    if data is not None:
        o.visible[...] = data
    elif rng is not None:
        v = rng.integers(0, 1 << bit_depth, size=(height, width))
        if smooth:
            v = v.astype(np.float64)
            for _ in range(3):
                v = (v + np.roll(v, 1, 0) + np.roll(v, 1, 1) + np.roll(v, -1, 0) + np.roll(v, -1, 1)) / 5
            v = np.clip(np.rint(v), 0, (1 << bit_depth) - 1)
        o.visible[...] = v.astype(o.dtype)
    if pad:
        o.pad()
    d = r.Plane(ctx, width, height, xdec, ydec, xpad, ypad, px_bytes).upload(o.data)
    return o, d


def oracle_full_search(r1o, o_org, o_ref, job, init=None):
    """job: one element of r.FS_JOB -> ((row, col), cost) or None on bounds error"""
    L = r1o.lib()
    bw, bh = L.r1o_block_width(int(job["bsize"])), L.r1o_block_height(int(job["bsize"]))
    best = r1o.Mv(0, 0)
    cost = C.c_uint64(U64_MAX)
    if init is not None:
        best = r1o.Mv(int(init[0][0]), int(init[0][1]))
        cost = C.c_uint64(int(init[1]))
    pmv = (r1o.Mv * 2)(r1o.Mv(int(job["pmv"][0]["row"]), int(job["pmv"][0]["col"])),
                        r1o.Mv(int(job["pmv"][1]["row"]), int(job["pmv"][1]["col"])))
    rc = L.r1o_full_search(o_org.origin_ptr(), C.byref(o_org.cfg), o_ref.origin_ptr(),
                           C.byref(o_ref.cfg), int(job["x_lo"]), int(job["x_hi"]),
                           int(job["y_lo"]), int(job["y_hi"]), bw, bh, int(job["po_x"]),
                           int(job["po_y"]), int(job["step"]), int(job["lambda_"]), C.byref(pmv),
                           int(job["allow_high_precision_mv"]), o_org.px_bytes, C.byref(best),
                           C.byref(cost))
    if rc != 0:
        return None
    return (best.row, best.col), cost.value


def full_pixel_window(r1o, w_in_b, h_in_b, po, bw, bh, cmv, rng_x, rng_y):
    """FullSearch::full_pixel_me window (src/me.rs:533-546) with a free range"""
    out = (C.c_int64 * 4)()
    r1o.lib().r1o_mv_range(w_in_b, h_in_b, po[0] // 4, po[1] // 4, bw, bh, C.byref(out))
    mvx_min, mvx_max, mvy_min, mvy_max = out

    def tdiv(a, b):  # Rust `/` on isize truncates toward zero
        q = abs(a) // b
        return q if a >= 0 else -q

    def clamp(v, lo, hi):
        return min(max(v, lo), hi)
    cx, cy = tdiv(cmv[1], 8), tdiv(cmv[0], 8)
    x_lo = po[0] + clamp(-rng_x + cx, tdiv(mvx_min, 8), tdiv(mvx_max, 8))
    x_hi = po[0] + clamp(rng_x + cx, tdiv(mvx_min, 8), tdiv(mvx_max, 8))
    y_lo = po[1] + clamp(-rng_y + cy, tdiv(mvy_min, 8), tdiv(mvy_max, 8))
    y_hi = po[1] + clamp(rng_y + cy, tdiv(mvy_min, 8), tdiv(mvy_max, 8))
    return x_lo, x_hi, y_lo, y_hi, tuple(out)


def oracle_subpel_cost(r1o, o_org, o_ref, job, mv, bit_depth):
    L = r1o.lib()
    bw, bh = L.r1o_block_width(int(job["bsize"])), L.r1o_block_height(int(job["bsize"]))
    pmv = (r1o.Mv * 2)(r1o.Mv(int(job["pmv"][0]["row"]), int(job["pmv"][0]["col"])),
                        r1o.Mv(int(job["pmv"][1]["row"]), int(job["pmv"][1]["col"])))
    rng = (C.c_int64 * 4)(int(job["mvx_min"]), int(job["mvx_max"]), int(job["mvy_min"]),
                          int(job["mvy_max"]))
    return L.r1o_subpel_cost(o_org.origin_ptr(), C.byref(o_org.cfg), o_ref.origin_ptr(),
                             C.byref(o_ref.cfg), int(job["po_x"]), int(job["po_y"]), bw, bh,
                             r1o.Mv(int(mv[0]), int(mv[1])), C.byref(pmv), int(job["lambda_"]),
                             int(job["use_satd"]), int(job["allow_high_precision_mv"]),
                             C.byref(rng), bit_depth, o_org.px_bytes)


def oracle_fullpel_satd_cost(r1o, o_org, o_ref, job, mv):
    """get_mv_rd_cost with tmp_plane_opt = None (src/me.rs:828-836), SATD"""
    L = r1o.lib()
    if not (job["mvx_min"] <= mv[1] <= job["mvx_max"] and job["mvy_min"] <= mv[0] <= job["mvy_max"]):
        return U64_MAX
    bw, bh = L.r1o_block_width(int(job["bsize"])), L.r1o_block_height(int(job["bsize"]))

    def tdiv(a, b):
        q = abs(a) // b
        return q if a >= 0 else -q
    d = r1o.satd(o_org, int(job["po_x"]), int(job["po_y"]), o_ref, int(job["po_x"]) + tdiv(mv[1], 8),
                 int(job["po_y"]) + tdiv(mv[0], 8), bw, bh)
    r1 = L.r1o_mv_rate(r1o.Mv(*mv), r1o.Mv(int(job["pmv"][0]["row"]), int(job["pmv"][0]["col"])),
                       int(job["allow_high_precision_mv"]))
    r2 = L.r1o_mv_rate(r1o.Mv(*mv), r1o.Mv(int(job["pmv"][1]["row"]), int(job["pmv"][1]["col"])),
                       int(job["allow_high_precision_mv"]))
    return 256 * d + min(r1, r2 + 1) * int(job["lambda_"])


def oracle_telescopic(r1o, o_org, o_ref, job, best_mv, lowest_cost, bit_depth):
    """motion_estimation tail for FullSearch (src/me.rs:232-271 + 858-941)"""
    best = tuple(best_mv)
    lowest = lowest_cost
    if int(job["use_satd"]):
        lowest = oracle_fullpel_satd_cost(r1o, o_org, o_ref, job, best)
    steps = [8, 4, 2] + ([1] if int(job["allow_high_precision_mv"]) else [])
    for step in steps:
        center = best
        for i in range(3):
            for j in range(3):
                if i == 1 and j == 1:
                    continue
                cand = (center[0] + step * (i - 1), center[1] + step * (j - 1))
                c = oracle_subpel_cost(r1o, o_org, o_ref, job, cand, bit_depth)
                if c < lowest:
                    lowest, best = c, cand
    return best, lowest
