"""ctypes binding of the CPU ORACLE (oracle/libr1o.so).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  The product package
(rav1e_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libr1o.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    if force or not os.path.exists(_SO) or any(
            os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "libr1o.so"], stdout=subprocess.DEVNULL)
    return _SO


class PlaneCfg(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "stride", "alloc_height", "width", "height", "xdec", "ydec", "xpad", "ypad",
        "xorigin", "yorigin")]


class QuantCtx(C.Structure):
    """r1o_quant_ctx = QuantizationContext, src/quantize.rs:111-121"""
    _fields_ = [("log_tx_scale", C.c_uint32), ("dc_quant", C.c_uint32), ("ac_quant", C.c_uint32),
                ("dc_offset", C.c_int32), ("ac_offset_eob", C.c_int32), ("ac_offset0", C.c_int32),
                ("ac_offset1", C.c_int32), ("dc_mul_add", C.c_uint32 * 3), ("ac_mul_add", C.c_uint32 * 3)]


class Mv(C.Structure):
    _fields_ = [("row", C.c_int16), ("col", C.c_int16)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        vp, i, i64, u32, u64, pd = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_ssize_t
        P = C.POINTER
        L.r1o_block_width.restype = i
        L.r1o_block_height.restype = i
        L.r1o_tx_width.restype = i
        L.r1o_tx_height.restype = i
        L.r1o_plane_cfg_new.argtypes = [P(PlaneCfg), i64, i64, i64, i64, i64, i64, i]
        L.r1o_plane_pad.argtypes = [vp, P(PlaneCfg), i64, i64, i]
        L.r1o_plane_downsample.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i]
        L.r1o_plane_downsample.restype = i
        for f in (L.r1o_sad, L.r1o_satd, L.r1o_satd_i16):
            f.argtypes = [vp, pd, vp, pd, i, i, i]
            f.restype = u32
        L.r1o_put_8tap.argtypes = [vp, pd, vp, pd, i, i, i, i, i, i, i, i]
        L.r1o_prep_8tap.argtypes = [vp, vp, pd, i, i, i, i, i, i, i, i]
        L.r1o_mc_avg.argtypes = [vp, pd, vp, vp, i, i, i, i]
        L.r1o_mv_params.argtypes = [P(PlaneCfg), i64, i64, i, i, P(i), P(i), P(i64), P(i64)]
        L.r1o_fwd_txfm1d.argtypes = [i, i, vp, vp]
        L.r1o_fwd_txfm1d.restype = i
        L.r1o_fht.argtypes = [vp, vp, i, i, i]
        L.r1o_fht.restype = i
        L.r1o_fht_valid.argtypes = [i, i]
        L.r1o_fht_valid.restype = i
        L.r1o_diff.argtypes = [vp, vp, pd, vp, pd, i, i, i]
        L.r1o_mv_rate.argtypes = [Mv, Mv, i]
        L.r1o_mv_rate.restype = u32
        L.r1o_mv_range.argtypes = [i64, i64, i64, i64, i, i, P(i64 * 4)]
        L.r1o_full_search.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i64, i64, i, i,
                                      i64, i64, i, u32, P(Mv * 2), i, i, P(Mv), P(u64)]
        L.r1o_full_search.restype = i
        L.r1o_subpel_cost.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i, i, Mv,
                                      P(Mv * 2), u32, i, i, P(i64 * 4), i, i]
        L.r1o_subpel_cost.restype = u64
        L.r1o_num_threads.restype = i
        L.r1o_bench_dist.argtypes = [i, vp, pd, vp, pd, i, i, i, i64]
        L.r1o_bench_dist.restype = C.c_double
        L.r1o_dist_batch.argtypes = [i, vp, P(PlaneCfg), vp, P(PlaneCfg), vp, i64, i, vp]
        L.r1o_full_search_batch.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), vp, i64, i, vp, vp]
        L.r1o_full_search_batch.restype = i
        L.r1o_mc_batch.argtypes = [i, vp, P(PlaneCfg), vp, vp, vp, P(PlaneCfg), vp, i64, i, i]
        L.r1o_avg_batch.argtypes = [vp, P(PlaneCfg), vp, vp, vp, vp, i64, i, i]
        L.r1o_inv_txfm1d.argtypes = [i, i, vp, vp, i]
        L.r1o_inv_txfm1d.restype = i
        L.r1o_inv_txfm_valid.argtypes = [i, i]
        L.r1o_inv_txfm_valid.restype = i
        L.r1o_inv_txfm2d.argtypes = [vp, vp, i, i, i]
        L.r1o_inv_txfm2d.restype = i
        L.r1o_inv_txfm_add_batch.argtypes = [vp, vp, i64, i, vp, i64, i, i, i]
        L.r1o_inv_txfm_add_batch.restype = i
        L.r1o_fht_batch.argtypes = [vp, vp, i, i, i, i64]
        L.r1o_fht_batch.restype = i
        L.r1o_frame_me_layout.argtypes = [i, i, u32, P(i64 * 22), P(C.c_int32 * 22), P(C.c_int32 * 22)]
        L.r1o_frame_me_layout.restype = i64
        L.r1o_frame_me.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i, i, i, i, u32, u32, i, i, i, i, i,
                                   vp, i64, vp, vp]
        L.r1o_frame_me.restype = i64
        L.r1o_fullpel_cost.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i, i, Mv, P(Mv * 2), u32, i, i,
                                       P(i64 * 4), i]
        L.r1o_fullpel_cost.restype = u64
        L.r1o_diamond_search.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i, i, vp, i, P(Mv * 2), u32, i, i,
                                         P(i64 * 4), i, i, i, P(Mv), P(u64)]
        L.r1o_diamond_search.restype = i
        dbl = C.c_double
        L.r1o_estimate_motion_ss4.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i64, i64, i, i, i, dbl, i, i, P(Mv)]
        L.r1o_estimate_motion_ss4.restype = i
        for f in (L.r1o_estimate_motion_ss2, L.r1o_estimate_motion):
            f.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), i64, i64, i64, i64, i, i, vp, vp, dbl, i, i, P(Mv)]
            f.restype = i
        L.r1o_distortion_bias.argtypes = [vp, i64, i64, i64, i64, i64, i64, i64]
        L.r1o_distortion_bias.restype = dbl
        L.r1o_sse_wxh.argtypes = [vp, i64, i64, i64, vp, i64, i64, i64, i64, i64, i, i, i, vp, i64, i64, i64]
        L.r1o_sse_wxh.restype = u64
        L.r1o_cdef_dist_wxh.argtypes = [vp, i64, i64, i64, vp, i64, i64, i64, i64, i64, i, i, vp, i64, i64, i64]
        L.r1o_cdef_dist_wxh.restype = u64
        L.r1o_tx_dist.argtypes = [vp, vp, i]
        L.r1o_tx_dist.restype = u64
        L.r1o_log_tx_scale.argtypes = [i]
        L.r1o_log_tx_scale.restype = i
        L.r1o_divu_gen.argtypes = [u32, P(u32 * 3)]
        L.r1o_divu_pair.argtypes = [C.c_int32, P(u32 * 3)]
        L.r1o_divu_pair.restype = C.c_int32
        L.r1o_quant_ctx_update.argtypes = [P(QuantCtx), i, i, u32, u32]
        L.r1o_quantize_batch.argtypes = [P(QuantCtx), i, vp, vp, vp, vp, i64]
        L.r1o_dequantize.argtypes = [u32, u32, i, vp, vp, i]
        L.r1o_importance_propagate.argtypes = [vp, P(PlaneCfg), vp, P(PlaneCfg), vp, i, vp, vp, i, i, i, i, i, vp]
        L.r1o_importance_finalize.argtypes = [vp, vp, i64]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OPlane:
    """Host plane with the reference's geometry (src/frame/plane.rs:215-244)."""

    def __init__(self, width, height, xdec=0, ydec=0, xpad=0, ypad=0, px_bytes=1, fill=128):
        self.cfg = PlaneCfg()
        lib().r1o_plane_cfg_new(C.byref(self.cfg), width, height, xdec, ydec, xpad, ypad, px_bytes)
        self.px_bytes = px_bytes
        self.dtype = np.uint8 if px_bytes == 1 else np.uint16
        self.data = np.full((self.cfg.alloc_height, self.cfg.stride), fill, dtype=self.dtype)

    @property
    def visible(self):
        c = self.cfg
        return self.data[c.yorigin:c.yorigin + c.height, c.xorigin:c.xorigin + c.width]

    def pad(self, w=None, h=None):
        c = self.cfg
        lib().r1o_plane_pad(_p(self.data), C.byref(self.cfg),
                            (c.width << c.xdec) if w is None else w,
                            (c.height << c.ydec) if h is None else h, self.px_bytes)

    def origin_ptr(self):
        """address of plane coordinate (0,0)"""
        c = self.cfg
        return self.data.ctypes.data + (c.yorigin * c.stride + c.xorigin) * self.px_bytes

    def ptr(self, x, y):
        return self.origin_ptr() + (y * self.cfg.stride + x) * self.px_bytes


def sad(org: OPlane, ox, oy, ref: OPlane, rx, ry, w, h):
    return lib().r1o_sad(org.ptr(ox, oy), org.cfg.stride, ref.ptr(rx, ry), ref.cfg.stride, w, h,
                         org.px_bytes)


def satd(org: OPlane, ox, oy, ref: OPlane, rx, ry, w, h):
    return lib().r1o_satd(org.ptr(ox, oy), org.cfg.stride, ref.ptr(rx, ry), ref.cfg.stride, w, h,
                          org.px_bytes)


def satd_i16(org: OPlane, ox, oy, ref: OPlane, rx, ry, w, h):
    """SATD as the fork's generated kernels compute it: i16 lanes (build/kernel/gen/dist.rs:171-350)"""
    return lib().r1o_satd_i16(org.ptr(ox, oy), org.cfg.stride, ref.ptr(rx, ry), ref.cfg.stride, w, h,
                              org.px_bytes)


def diamond_search(org: OPlane, ref: OPlane, po, blk_w, blk_h, predictors, pmv, lambda_, use_satd, allow_hp, mv_range,
                   subpixel, bit_depth):
    """diamond_me_search + get_best_predictor, src/me.rs:655-785 -> ((row, col), cost, iterations)"""
    pred = np.ascontiguousarray(np.asarray(predictors, dtype=np.int16).reshape(-1, 2))  # (row, col)
    pm = (Mv * 2)(Mv(*pmv[0]), Mv(*pmv[1]))
    rng = (C.c_int64 * 4)(*mv_range)
    mv, cost = Mv(0, 0), C.c_uint64(0)
    it = lib().r1o_diamond_search(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(), C.byref(ref.cfg), po[0], po[1],
                                  blk_w, blk_h, _p(pred), pred.shape[0], C.byref(pm), lambda_, int(use_satd),
                                  int(allow_hp), C.byref(rng), int(subpixel), bit_depth, org.px_bytes, C.byref(mv),
                                  C.byref(cost))
    return (mv.row, mv.col), cost.value, it


def estimate_motion_ss4(org_q: OPlane, ref_q: OPlane, w_in_b, h_in_b, bo, blk_w, blk_h, me_range_scale, me_lambda, allow_hp):
    """estimate_motion_ss4, src/me.rs:1023-1075 -> (row, col)"""
    mv = Mv(0, 0)
    rc = lib().r1o_estimate_motion_ss4(org_q.origin_ptr(), C.byref(org_q.cfg), ref_q.origin_ptr(), C.byref(ref_q.cfg), w_in_b,
                                       h_in_b, bo[0], bo[1], blk_w, blk_h, me_range_scale, float(me_lambda), int(allow_hp),
                                       org_q.px_bytes, C.byref(mv))
    assert rc == 0
    return mv.row, mv.col


def estimate_motion_pred(org: OPlane, ref: OPlane, w_in_b, h_in_b, bo, blk_w, blk_h, preds, valid, me_lambda, allow_hp, ss2):
    """estimate_motion_ss2 (ss2=True, 3 pmvs, half-res planes, src/me.rs:280-333,597-650) or estimate_motion with
    FullSearch::full_pixel_me (7 cmvs, :335-385,522-565) -> (row, col)"""
    slots = 3 if ss2 else 7
    p = np.ascontiguousarray(np.asarray(preds, dtype=np.int16).reshape(slots, 2))
    v = np.ascontiguousarray(np.asarray(valid, dtype=np.uint8).reshape(slots))
    mv = Mv(0, 0)
    fn = lib().r1o_estimate_motion_ss2 if ss2 else lib().r1o_estimate_motion
    rc = fn(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(), C.byref(ref.cfg), w_in_b, h_in_b, bo[0], bo[1], blk_w, blk_h,
            _p(p), _p(v), float(me_lambda), int(allow_hp), org.px_bytes, C.byref(mv))
    assert rc == 0
    return mv.row, mv.col


def _tdiv(a, b):
    """Rust integer `/`: truncation toward zero"""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b > 0) else -q


def _i16(v):
    return ((v + 32768) & 0xFFFF) - 32768


def save_block_motion(field, bo, blk_w, blk_h, mv):
    """save_block_motion, src/encoder.rs:1432-1444; field: int16 array (rows, cols, 2) as (row, col)"""
    rows, cols = field.shape[:2]
    for y in range(bo[1], min(bo[1] + blk_h // 4, rows)):
        for x in range(bo[0], min(bo[0] + blk_w // 4, cols)):
            field[y, x] = mv


def get_subset_predictors(bo, cmvs, field, prev_field):
    """get_subset_predictors, src/me.rs:82-174 (tile == frame) -> list of (row, col)"""
    def fullpel(m):
        return (_tdiv(m[0], 8) * 8, _tdiv(m[1], 8) * 8)

    rows, cols = field.shape[:2]
    x, y = bo
    preds = [(0, 0)] + [fullpel(m) for m in cmvs]
    med = []
    if x > 0:
        left = tuple(int(v) for v in field[y, x - 1])
        med.append(left)
        if left != (0, 0):
            preds.append(left)
    if y > 0:
        top = tuple(int(v) for v in field[y - 1, x])
        med.append(top)
        if top != (0, 0):
            preds.append(top)
        if x < cols - 1:
            tr = tuple(int(v) for v in field[y - 1, x + 1])
            med.append(tr)
            if tr != (0, 0):
                preds.append(tr)
    if med:
        sr, sc = 0, 0
        for m in med:
            sr, sc = _i16(sr + m[0]), _i16(sc + m[1])
        q = fullpel((_tdiv(sr, len(med)), _tdiv(sc, len(med))))
        if q != (0, 0):
            preds.append(q)
    if prev_field is not None:
        cand = []
        if x > 0:
            cand.append(prev_field[y, x - 1])
        if y > 0:
            cand.append(prev_field[y - 1, x])
        if x < cols - 1:
            cand.append(prev_field[y, x + 1])
        if y < rows - 1:
            cand.append(prev_field[y + 1, x])
        cand.append(prev_field[y, x])
        for m in cand:
            m = (int(m[0]), int(m[1]))
            if m != (0, 0):
                preds.append(m)
    return preds


def _imp(block_importances):
    if block_importances is None:
        return None, None
    a = np.ascontiguousarray(block_importances, dtype=np.float32)
    return a, _p(a)


def sse_wxh(src1: OPlane, x1, y1, src2: OPlane, x2, y2, w, h, block_importances=None, w_in_imp_b=0, w_in_b=0,
            h_in_b=0):
    """sse_wxh, src/rdo.rs:286-335 (bias from block_importances; None -> 1.0)"""
    _keep, ip = _imp(block_importances)
    return lib().r1o_sse_wxh(src1.origin_ptr(), src1.cfg.stride, x1, y1, src2.origin_ptr(), src2.cfg.stride, x2, y2,
                             w, h, src1.cfg.xdec, src1.cfg.ydec, src1.px_bytes, ip, w_in_imp_b, w_in_b, h_in_b)


def cdef_dist_wxh(src1: OPlane, x1, y1, src2: OPlane, x2, y2, w, h, bit_depth, block_importances=None, w_in_imp_b=0,
                  w_in_b=0, h_in_b=0):
    """cdef_dist_wxh, src/rdo.rs:256-283"""
    _keep, ip = _imp(block_importances)
    return lib().r1o_cdef_dist_wxh(src1.origin_ptr(), src1.cfg.stride, x1, y1, src2.origin_ptr(), src2.cfg.stride,
                                   x2, y2, w, h, bit_depth, src1.px_bytes, ip, w_in_imp_b, w_in_b, h_in_b)


def tx_dist(coeffs, rcoeffs, tx_size):
    """tx-domain distortion of one block, src/encoder.rs:1210-1224"""
    c = np.ascontiguousarray(coeffs, dtype=np.int32)
    r = np.ascontiguousarray(rcoeffs, dtype=np.int32)
    return lib().r1o_tx_dist(_p(c), _p(r), tx_size)


def quant_ctx(tx_size, is_intra, dc_quant, ac_quant):
    """QuantizationContext::update, src/quantize.rs:212-255, from the dc_q / ac_q values"""
    qc = QuantCtx()
    lib().r1o_quant_ctx_update(C.byref(qc), tx_size, int(bool(is_intra)), dc_quant, ac_quant)
    return qc


def quantize_batch(qc, tx_size, scan, coeffs, want_rcoeffs=False):
    """QuantizationContext::quantize (src/quantize.rs:257-316) of N blocks (N, H*W) -> (N, coded);
    with want_rcoeffs also dequantize (:318-334) of the result"""
    L = lib()
    w, h = L.r1o_tx_width(tx_size), L.r1o_tx_height(tx_size)
    coded = min(w, 32) * min(h, 32)
    c = np.ascontiguousarray(coeffs, dtype=np.int32)
    sc = np.ascontiguousarray(scan, dtype=np.uint16)
    n = c.size // (w * h)
    assert sc.size == coded and n * w * h == c.size
    q = np.empty((n, coded), dtype=np.int32)
    r = np.empty((n, coded), dtype=np.int32) if want_rcoeffs else None
    L.r1o_quantize_batch(C.byref(qc), tx_size, _p(sc), _p(c), _p(q), _p(r) if want_rcoeffs else None, n)
    return (q, r) if want_rcoeffs else q


def dequantize(qcoeffs, dc_quant, ac_quant, log_tx_scale):
    """dequantize, src/quantize.rs:318-334, of one block"""
    q = np.ascontiguousarray(qcoeffs, dtype=np.int32)
    r = np.empty_like(q)
    lib().r1o_dequantize(dc_quant, ac_quant, log_tx_scale, _p(q), _p(r), q.size)
    return r


def divu_gen(d):
    out = (C.c_uint32 * 3)()
    lib().r1o_divu_gen(d, C.byref(out))
    return out


def importance_propagate(frame, ref, mvs, intra_costs, importances, n_refs, bit_depth, ref_importances):
    """one (frame, reference) pair of compute_block_importances, src/api/internal.rs:897-1047; mvs int16 (rows, cols, 2)
    = {row, col}; ref_importances float32 (h, w) is updated in place"""
    mv = np.ascontiguousarray(mvs, dtype=np.int16)
    ic = np.ascontiguousarray(intra_costs, dtype=np.uint32)
    im = np.ascontiguousarray(importances, dtype=np.float32)
    assert ref_importances.dtype == np.float32 and ref_importances.flags.c_contiguous
    h, w = ic.shape
    lib().r1o_importance_propagate(frame.origin_ptr(), C.byref(frame.cfg), ref.origin_ptr(), C.byref(ref.cfg), _p(mv),
                                   mv.shape[1], _p(ic), _p(im), n_refs, w, h, bit_depth, frame.px_bytes, _p(ref_importances))
    return ref_importances


def importance_finalize(intra_costs, importances):
    """src/api/internal.rs:1054-1071, in place"""
    ic = np.ascontiguousarray(intra_costs, dtype=np.uint32)
    assert importances.dtype == np.float32 and importances.flags.c_contiguous
    lib().r1o_importance_finalize(_p(ic), _p(importances), ic.size)
    return importances


def lookahead_intra_costs(plane: OPlane, w_in_imp_b, h_in_imp_b, bit_depth):
    """compute_lookahead_intra_costs, src/api/internal.rs:680-765: tile_rect = the block itself -> PredictionVariant::NONE
    -> pred_dc_128 (src/predict.rs:176-184,623-633), then get_satd of the 8x8 block"""
    flat = OPlane(8, 8, 0, 0, 0, 0, plane.px_bytes)
    flat.data[...] = 128 << (bit_depth - 8)
    out = np.zeros((h_in_imp_b, w_in_imp_b), dtype=np.uint32)
    for y in range(h_in_imp_b):
        for x in range(w_in_imp_b):
            out[y, x] = satd(plane, x * 8, y * 8, flat, 0, 0, 8, 8)
    return out


def fht(residual, tx_size, tx_type, bit_depth):
    residual = np.ascontiguousarray(residual, dtype=np.int16)
    out = np.empty(residual.size, dtype=np.int32)
    rc = lib().r1o_fht(_p(residual), _p(out), tx_size, tx_type, bit_depth)
    if rc != 0:
        raise ValueError("invalid (tx_size, tx_type, bit_depth)")
    return out.reshape(residual.shape)


def fht_batch(residual, tx_size, tx_type, bit_depth):
    residual = np.ascontiguousarray(residual, dtype=np.int16)
    area = lib().r1o_tx_width(tx_size) * lib().r1o_tx_height(tx_size)
    n = residual.size // area
    out = np.empty(residual.size, dtype=np.int32)
    rc = lib().r1o_fht_batch(_p(residual), _p(out), tx_size, tx_type, bit_depth, n)
    if rc != 0:
        raise ValueError("invalid (tx_size, tx_type)")
    return out


def inv_txfm1d(type1d, vec, rng):
    """src/transform/inverse.rs 1-D functions; type1d: 0 DCT, 1 ADST, 2 FLIPADST, 3 identity"""
    vec = np.ascontiguousarray(vec, dtype=np.int32)
    out = np.empty_like(vec)
    if lib().r1o_inv_txfm1d(type1d, vec.size, _p(vec), _p(out), rng) != 0:
        raise ValueError("unimplemented (type, n)")
    return out


def inv_txfm_valid(tx_size, tx_type):
    return bool(lib().r1o_inv_txfm_valid(tx_size, tx_type))


def coded_area(tx_size):
    return min(lib().r1o_tx_width(tx_size), 32) * min(lib().r1o_tx_height(tx_size), 32)


def inv_txfm2d(coeffs, tx_size, tx_type, bit_depth):
    """inv_txfm2d, inverse.rs:2061-2107 -> (H, W) residual (the reference's column-major array, transposed)"""
    w, h = lib().r1o_tx_width(tx_size), lib().r1o_tx_height(tx_size)
    coeffs = np.ascontiguousarray(coeffs, dtype=np.int32)
    assert coeffs.size == coded_area(tx_size)
    out = np.empty(w * h, dtype=np.int32)
    if lib().r1o_inv_txfm2d(_p(coeffs), _p(out), tx_size, tx_type, bit_depth) != 0:
        raise ValueError("invalid (tx_size, tx_type, bit_depth)")
    return out.reshape(w, h).T.copy()


def inv_txfm_add_batch(coeffs, plane, xy, tx_size, tx_type, bit_depth):
    """inv_txfm2d_add, inverse.rs:1939-2028, over n non-overlapping blocks; `plane` (2-D u8 / u16 array) is
    updated in place, block i at column xy[i,0], row xy[i,1]"""
    coeffs = np.ascontiguousarray(coeffs, dtype=np.int32)
    xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
    assert plane.flags.c_contiguous and plane.dtype in (np.uint8, np.uint16)
    assert coeffs.size == xy.shape[0] * coded_area(tx_size)
    rc = lib().r1o_inv_txfm_add_batch(_p(coeffs), _p(plane), plane.shape[1], plane.dtype.itemsize, _p(xy),
                                      xy.shape[0], tx_size, tx_type, bit_depth)
    if rc != 0:
        raise ValueError("invalid (tx_size, tx_type, bit_depth)")
    return plane


def txfm1d(type1d, vec):
    vec = np.ascontiguousarray(vec, dtype=np.int32)
    out = np.empty_like(vec)
    rc = lib().r1o_fwd_txfm1d(type1d, vec.size, _p(vec), _p(out))
    if rc != 0:
        raise ValueError("unimplemented (type, n)")
    return out


def frame_me(org: OPlane, ref: OPlane, range_x, range_y, lambda_, bsize_mask, refine, use_satd,
             allow_hp, bit_depth, idx=None):
    """motion_estimation (src/me.rs:193-271) for every block of the frame (or the frame-layout
    slots `idx`) -> (mv int16 (n, 2), cost uint64 (n,)); untouched slots hold cost 0 / mv 0"""
    L = lib()
    w, h = int(org.cfg.width), int(org.cfg.height)
    off, cols, rows = (C.c_int64 * 22)(), (C.c_int32 * 22)(), (C.c_int32 * 22)()
    total = L.r1o_frame_me_layout(w, h, bsize_mask, C.byref(off), C.byref(cols), C.byref(rows))
    mv = np.zeros((total, 2), dtype=np.int16)
    cost = np.zeros(total, dtype=np.uint64)
    ip, n_idx = None, 0
    if idx is not None:
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        ip, n_idx = _p(idx), idx.size
    L.r1o_frame_me(org.origin_ptr(), C.byref(org.cfg), ref.origin_ptr(), C.byref(ref.cfg), w, h,
                   range_x, range_y, lambda_, bsize_mask, int(refine), int(use_satd), int(allow_hp),
                   bit_depth, org.px_bytes, ip, n_idx, _p(mv), _p(cost))
    return mv, cost
