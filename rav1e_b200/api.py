"""Host-side mirror of the reference's interface for the hot path, on top of the
C ABI (include/rav1e_cuda.h).  Names and argument meaning follow the reference:

    Plane / PlaneConfig          src/frame/plane.rs:22-47, 154, 215-244
    get_sad / get_satd           src/dist.rs:48, 121
    put_8tap / prep_8tap / mc_avg  src/mc.rs:410, 520, 617
    forward_transform            src/transform/mod.rs:556
    full_search                  src/me.rs:943
    get_mv_range / get_mv_rate   src/me.rs:64, 1005

plus the batched entry points (job arrays are numpy structured arrays whose
dtypes mirror the C structs).  Everything here runs on the GPU through
librav1e_cuda.so; there is no CPU path.  Errors of the C ABI surface as
Rav1eCudaError (the reference `assert!`s / panics in the same situations).
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import Rav1eCudaError  # noqa: F401  (re-export)

# numpy mirrors of the job / result structs of include/rav1e_cuda.h
MV = np.dtype([("row", "<i2"), ("col", "<i2")])
DIST_JOB = np.dtype([("org_x", "<i4"), ("org_y", "<i4"), ("ref_x", "<i4"), ("ref_y", "<i4"),
                     ("bsize", "<i4")])
MC_JOB = np.dtype([("dst_x", "<i4"), ("dst_y", "<i4"), ("src_x", "<i4"), ("src_y", "<i4"),
                   ("w", "<i2"), ("h", "<i2"), ("col_frac", "i1"), ("row_frac", "i1"),
                   ("mode_x", "i1"), ("mode_y", "i1"), ("tmp_off", "<i4")])
AVG_JOB = np.dtype([("dst_x", "<i4"), ("dst_y", "<i4"), ("w", "<i2"), ("h", "<i2"),
                    ("tmp_off", "<i4")])
TX_JOB = np.dtype([("in_x", "<i4"), ("in_y", "<i4"), ("pred_x", "<i4"), ("pred_y", "<i4")])
ITX_JOB = np.dtype([("x", "<i4"), ("y", "<i4")])
BLOCK_JOB = np.dtype([("bo_x", "<i4"), ("bo_y", "<i4"), ("bsize", "<i4")])
RDO_JOB = np.dtype([("src_x", "<i4"), ("src_y", "<i4"), ("rec_x", "<i4"), ("rec_y", "<i4"),
                    ("w", "<i4"), ("h", "<i4")])
ME_RESULT = np.dtype([("mv", MV), ("reserved", "<u4"), ("cost", "<u8")])
FS_JOB = np.dtype([("po_x", "<i4"), ("po_y", "<i4"), ("bsize", "<i4"), ("x_lo", "<i4"),
                   ("x_hi", "<i4"), ("y_lo", "<i4"), ("y_hi", "<i4"), ("step", "<i4"),
                   ("lambda_", "<u4"), ("pmv", MV, (2,)), ("allow_high_precision_mv", "<i4"),
                   ("has_init", "<i4")])
SB_JOB = np.dtype([("po_x", "<i4"), ("po_y", "<i4"), ("bsize_mask", "<u4"), ("cmv", MV),
                   ("pmv", MV, (2,))])
SUBPEL_JOB = np.dtype([("po_x", "<i4"), ("po_y", "<i4"), ("bsize", "<i4"), ("lambda_", "<u4"),
                       ("pmv", MV, (2,)), ("mvx_min", "<i4"), ("mvx_max", "<i4"),
                       ("mvy_min", "<i4"), ("mvy_max", "<i4"), ("use_satd", "<i4"),
                       ("allow_high_precision_mv", "<i4")])
STRUCT_PAIRS = [(DIST_JOB, capi.DistJob), (MC_JOB, capi.McJob), (AVG_JOB, capi.AvgJob),
                (TX_JOB, capi.TxJob), (ITX_JOB, capi.ItxJob), (RDO_JOB, capi.RdoJob), (BLOCK_JOB, capi.BlockJob), (ME_RESULT, capi.MeResult), (FS_JOB, capi.FsJob),
                (SB_JOB, capi.SbJob), (SUBPEL_JOB, capi.SubpelJob), (MV, capi.Mv)]

BLOCK = {n: i for i, n in enumerate(capi.BLOCK_SIZES)}
TX_SIZE = {n: i for i, n in enumerate(capi.TX_SIZES)}
TX_TYPE = {n: i for i, n in enumerate(capi.TX_TYPES)}
FILTER = {n: i for i, n in enumerate(capi.FILTER_MODES)}
SB_RESULTS = capi.R1C_SB_RESULTS


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _arr(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


class Context:
    """One device + stream + scratch (r1c_ctx).  One per worker thread / rank."""

    def __init__(self, device=0):
        self.lib = capi.load()
        self.h = self.lib.r1c_create(device)
        if not self.h:
            raise Rav1eCudaError(capi.R1C_ERR_CUDA, "no usable CUDA device %d (rav1e_b200 has no "
                                 "CPU fallback)" % device)
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.r1c_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != 0:
            raise Rav1eCudaError(rc, self.lib.r1c_last_error(self.h).decode())

    def sync(self):
        self.check(self.lib.r1c_sync(self.h))

    @property
    def stream(self):
        return self.lib.r1c_stream(self.h)

    @property
    def launches(self):
        return self.lib.r1c_launch_count(self.h)


class Plane:
    """Device-resident Plane<T> with the reference's geometry (Plane::new)."""

    def __init__(self, ctx, width, height, xdec=0, ydec=0, xpad=0, ypad=0, px_bytes=1,
                 cfg=None, device_ptr=None):
        self.ctx = ctx
        self.px_bytes = px_bytes
        self.dtype = np.uint8 if px_bytes == 1 else np.uint16
        self.cfg = capi.PlaneCfg()
        if cfg is not None:
            C.memmove(C.byref(self.cfg), C.byref(cfg), C.sizeof(capi.PlaneCfg))
        else:
            rc = ctx.lib.r1c_plane_cfg_new(C.byref(self.cfg), width, height, xdec, ydec, xpad,
                                           ypad, px_bytes)
            if rc:
                raise Rav1eCudaError(rc, "bad plane geometry")
        h = C.c_void_p()
        if device_ptr is None:
            ctx.check(ctx.lib.r1c_plane_create(ctx.h, C.byref(self.cfg), px_bytes, C.byref(h)))
        else:
            ctx.check(ctx.lib.r1c_plane_wrap(ctx.h, device_ptr, C.byref(self.cfg), px_bytes,
                                             C.byref(h)))
        self.h = h

    @property
    def shape(self):
        return (self.cfg.alloc_height, self.cfg.stride)

    def upload(self, host):
        """host: the WHOLE allocation, shape (alloc_height, stride)."""
        host = _arr(host, self.dtype)
        assert host.shape == self.shape, (host.shape, self.shape)
        self.ctx.check(self.ctx.lib.r1c_plane_upload(self.ctx.h, self.h, _p(host)))
        return self

    def download(self):
        out = np.empty(self.shape, dtype=self.dtype)
        self.ctx.check(self.ctx.lib.r1c_plane_download(self.ctx.h, self.h, _p(out)))
        return out

    def pad(self, w, h):
        self.ctx.check(self.ctx.lib.r1c_plane_pad(self.ctx.h, self.h, w, h))

    def downsample_from(self, src):
        self.ctx.check(self.ctx.lib.r1c_plane_downsample(self.ctx.h, self.h, src.h))

    @property
    def device_ptr(self):
        return self.ctx.lib.r1c_plane_device_ptr(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.r1c_plane_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------ distortion --
def get_sad(ctx, plane_org, org_xy, plane_ref, ref_xy, bsize, bit_depth):
    out = C.c_uint32()
    ctx.check(ctx.lib.r1c_sad(ctx.h, plane_org.h, org_xy[0], org_xy[1], plane_ref.h, ref_xy[0],
                              ref_xy[1], bsize, bit_depth, C.byref(out)))
    return out.value


def get_satd(ctx, plane_org, org_xy, plane_ref, ref_xy, bsize, bit_depth):
    out = C.c_uint32()
    ctx.check(ctx.lib.r1c_satd(ctx.h, plane_org.h, org_xy[0], org_xy[1], plane_ref.h, ref_xy[0],
                               ref_xy[1], bsize, bit_depth, C.byref(out)))
    return out.value


def sad_batch(ctx, plane_org, plane_ref, jobs, bit_depth, satd=False):
    jobs = _arr(jobs, DIST_JOB)
    out = np.empty(jobs.size, dtype=np.uint32)
    fn = ctx.lib.r1c_satd_batch if satd else ctx.lib.r1c_sad_batch
    ctx.check(fn(ctx.h, plane_org.h, plane_ref.h, _p(jobs), jobs.size, bit_depth, _p(out)))
    return out


def satd_batch(ctx, plane_org, plane_ref, jobs, bit_depth, compat_i16_wrap=False):
    """get_satd over a batch.  compat_i16_wrap: the arithmetic of the kernels the fork generates (i16 lanes for every
    pixel type, build/kernel/gen/dist.rs:171-350) instead of get_satd_ref's i32 -- differs on high-amplitude 10- / 12-bit
    blocks only; for comparisons with a reference binary."""
    if compat_i16_wrap:
        jobs = _arr(jobs, DIST_JOB)
        out = np.empty(jobs.size, dtype=np.uint32)
        ctx.check(ctx.lib.r1c_satd_i16_batch(ctx.h, plane_org.h, plane_ref.h, _p(jobs), jobs.size, bit_depth, _p(out)))
        return out
    return sad_batch(ctx, plane_org, plane_ref, jobs, bit_depth, satd=True)


# ------------------------------------------------------- motion compensation --
def lookahead_intra_costs(ctx, plane, w_in_imp_b, h_in_imp_b, bit_depth):
    """compute_lookahead_intra_costs (src/api/internal.rs:680-765) -> uint32 (h_in_imp_b, w_in_imp_b)"""
    out = np.zeros((h_in_imp_b, w_in_imp_b), dtype=np.uint32)
    ctx.check(ctx.lib.r1c_lookahead_intra_costs(ctx.h, plane.h, w_in_imp_b, h_in_imp_b, bit_depth, _p(out)))
    return out


def mv_params(rec_cfg, po, mv):
    """get_params of predict_inter: -> (row_frac, col_frac, src_x, src_y)"""
    lib = capi.load()
    rf, cf, sx, sy = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
    lib.r1c_mv_params(C.byref(rec_cfg), po[0], po[1], mv[0], mv[1], C.byref(rf), C.byref(cf),
                      C.byref(sx), C.byref(sy))
    return rf.value, cf.value, sx.value, sy.value


def put_8tap_batch(ctx, dst, src, jobs, bit_depth):
    jobs = _arr(jobs, MC_JOB)
    ctx.check(ctx.lib.r1c_put_8tap_batch(ctx.h, dst.h, src.h, _p(jobs), jobs.size, bit_depth))


def prep_8tap_batch(ctx, tmp, src, jobs, bit_depth):
    """tmp: int16 array written in place at jobs[i].tmp_off"""
    jobs = _arr(jobs, MC_JOB)
    assert tmp.dtype == np.int16 and tmp.flags.c_contiguous
    ctx.check(ctx.lib.r1c_prep_8tap_batch(ctx.h, _p(tmp), tmp.size, src.h, _p(jobs), jobs.size,
                                          bit_depth))


def mc_avg_batch(ctx, dst, tmp1, tmp2, jobs, bit_depth):
    jobs = _arr(jobs, AVG_JOB)
    tmp1 = _arr(tmp1, np.int16)
    tmp2 = _arr(tmp2, np.int16)
    assert tmp1.size == tmp2.size
    ctx.check(ctx.lib.r1c_mc_avg_batch(ctx.h, dst.h, _p(tmp1), _p(tmp2), tmp1.size, _p(jobs),
                                       jobs.size, bit_depth))


def put_8tap(ctx, dst, dst_xy, src, src_xy, width, height, col_frac, row_frac, mode_x, mode_y,
             bit_depth):
    j = np.zeros(1, dtype=MC_JOB)
    j[0] = (dst_xy[0], dst_xy[1], src_xy[0], src_xy[1], width, height, col_frac, row_frac,
            mode_x, mode_y, 0)
    put_8tap_batch(ctx, dst, src, j, bit_depth)


def prep_8tap(ctx, src, src_xy, width, height, col_frac, row_frac, mode_x, mode_y, bit_depth):
    j = np.zeros(1, dtype=MC_JOB)
    j[0] = (0, 0, src_xy[0], src_xy[1], width, height, col_frac, row_frac, mode_x, mode_y, 0)
    tmp = np.zeros(width * height, dtype=np.int16)
    prep_8tap_batch(ctx, tmp, src, j, bit_depth)
    return tmp.reshape(height, width)


def mc_avg(ctx, dst, dst_xy, tmp1, tmp2, width, height, bit_depth):
    j = np.zeros(1, dtype=AVG_JOB)
    j[0] = (dst_xy[0], dst_xy[1], width, height, 0)
    mc_avg_batch(ctx, dst, tmp1.reshape(-1), tmp2.reshape(-1), j, bit_depth)


# ------------------------------------------------------------ forward transform --
def fwd_txfm_valid(tx_size, tx_type):
    return bool(capi.load().r1c_fwd_txfm_valid(tx_size, tx_type))


def _bias_args(block_importances, w_in_imp_b, w_in_b, h_in_b):
    if block_importances is None:
        return None, None, None
    imp = _arr(block_importances, np.float32)
    bp = capi.BiasParams(int(w_in_imp_b), int(imp.size), int(w_in_b), int(h_in_b))
    return imp, bp, C.byref(bp)


def sse_wxh(ctx, src1, src2, jobs, bit_depth, block_importances=None, w_in_imp_b=0, w_in_b=0, h_in_b=0,
            cdef=False):
    """sse_wxh (src/rdo.rs:286-335) or, with cdef=True, cdef_dist_wxh (:256-283) of every RDO_JOB;
    the bias closure is compute_distortion_bias (:477-508) over block_importances (None: bias 1.0).
    Returns the u64 Distortion per job."""
    jobs = _arr(jobs, RDO_JOB)
    out = np.empty(jobs.size, dtype=np.uint64)
    imp, _keep, bp = _bias_args(block_importances, w_in_imp_b, w_in_b, h_in_b)
    fn = ctx.lib.r1c_cdef_dist_batch if cdef else ctx.lib.r1c_sse_wxh_batch
    ctx.check(fn(ctx.h, src1.h, src2.h, _p(jobs), jobs.size, bit_depth, _p(imp) if imp is not None else None,
                 bp, _p(out)))
    return out


def cdef_dist_wxh(ctx, src1, src2, jobs, bit_depth, block_importances=None, w_in_imp_b=0, w_in_b=0, h_in_b=0):
    return sse_wxh(ctx, src1, src2, jobs, bit_depth, block_importances, w_in_imp_b, w_in_b, h_in_b, cdef=True)


def tx_distortion(ctx, coeffs, rcoeffs, tx_size):
    """tx-domain distortion of encode_tx_block (src/encoder.rs:1210-1224); coeffs (N, H*W) int32,
    rcoeffs (N, min(H,32)*min(W,32)) int32 -> (N,) uint64"""
    coeffs, rcoeffs = _arr(coeffs, np.int32), _arr(rcoeffs, np.int32)
    w, h = ctx.lib.r1c_tx_width(tx_size), ctx.lib.r1c_tx_height(tx_size)
    n = coeffs.size // (w * h)
    assert n * w * h == coeffs.size and n * min(w, 32) * min(h, 32) == rcoeffs.size
    out = np.empty(n, dtype=np.uint64)
    ctx.check(ctx.lib.r1c_tx_dist_batch(ctx.h, _p(coeffs), _p(rcoeffs), tx_size, n, _p(out)))
    return out


def importance_propagate(ctx, frame, ref, mvs, intra_costs, importances, n_refs, bit_depth, ref_importances):
    """one (frame, reference) pair of compute_block_importances (src/api/internal.rs:897-1047).
    mvs: MV array (mv_rows, mv_cols) in MI units; intra_costs uint32 (h, w); importances float32 (h, w);
    ref_importances float32 (h, w), updated in place and returned"""
    mvs = _arr(mvs, MV)
    intra_costs, importances = _arr(intra_costs, np.uint32), _arr(importances, np.float32)
    assert ref_importances.dtype == np.float32 and ref_importances.flags.c_contiguous
    h, w = intra_costs.shape
    assert importances.shape == (h, w) and ref_importances.shape == (h, w) and mvs.ndim == 2
    ctx.check(ctx.lib.r1c_importance_propagate(ctx.h, frame.h, ref.h, _p(mvs), mvs.shape[1], mvs.shape[0], _p(intra_costs),
                                               _p(importances), n_refs, w, h, bit_depth, _p(ref_importances)))
    return ref_importances


def importance_finalize(ctx, intra_costs, importances):
    """the last loop of compute_block_importances (src/api/internal.rs:1054-1071), in place"""
    intra_costs = _arr(intra_costs, np.uint32)
    assert importances.dtype == np.float32 and importances.flags.c_contiguous and importances.size == intra_costs.size
    ctx.check(ctx.lib.r1c_importance_finalize(ctx.h, _p(intra_costs), _p(importances), intra_costs.size))
    return importances


class QuantizationContext:
    """QuantizationContext (src/quantize.rs:111-121, 212-316).  `update` takes the quantizer VALUES
    dc_q(qindex, dc_delta_q, bit_depth) / ac_q(..) -- the qindex tables (:41-61) stay with the caller."""

    def __init__(self):
        self.c = capi.QuantCtx()
        self.tx_size = None

    def update(self, lib, tx_size, is_intra, dc_quant, ac_quant):
        rc = lib.r1c_quant_ctx_update(C.byref(self.c), tx_size, int(bool(is_intra)), dc_quant, ac_quant)
        if rc:
            raise Rav1eCudaError(rc, "bad tx_size / quantizer")
        self.tx_size = tx_size
        return self

    def quantize(self, ctx, coeffs, scan):
        """coeffs (N, H*W) int32, scan = av1_scan_orders[tx_size][tx_type].scan (uint16, coded area)
        -> qcoeffs (N, min(H,32)*min(W,32)) int32"""
        coeffs, scan = _arr(coeffs, np.int32), _arr(scan, np.uint16)
        w, h = ctx.lib.r1c_tx_width(self.tx_size), ctx.lib.r1c_tx_height(self.tx_size)
        coded = min(w, 32) * min(h, 32)
        n = coeffs.size // (w * h)
        assert n * w * h == coeffs.size and scan.size == coded
        q = np.empty((n, coded), dtype=np.int32)
        ctx.check(ctx.lib.r1c_quantize_batch(ctx.h, C.byref(self.c), self.tx_size, _p(scan), _p(coeffs), _p(q), n))
        return q


def dequantize(ctx, qcoeffs, tx_size, dc_quant, ac_quant):
    """dequantize (src/quantize.rs:318-334) of (N, coded) int32 -> rcoeffs, same shape"""
    qcoeffs = _arr(qcoeffs, np.int32)
    w, h = ctx.lib.r1c_tx_width(tx_size), ctx.lib.r1c_tx_height(tx_size)
    coded = min(w, 32) * min(h, 32)
    n = qcoeffs.size // coded
    assert n * coded == qcoeffs.size
    r = np.empty(qcoeffs.shape, dtype=np.int32)
    ctx.check(ctx.lib.r1c_dequantize_batch(ctx.h, tx_size, dc_quant, ac_quant, _p(qcoeffs), _p(r), n))
    return r


def forward_transform(ctx, residual, tx_size, tx_type, bit_depth):
    """residual: int16 (H, W) or a batch (N, H, W) -> int32 coefficients, same shape."""
    residual = _arr(residual, np.int16)
    w = ctx.lib.r1c_tx_width(tx_size)
    h = ctx.lib.r1c_tx_height(tx_size)
    if w < 0:
        raise Rav1eCudaError(capi.R1C_ERR_ARG, "bad tx_size")
    n = residual.size // (w * h)
    assert n * w * h == residual.size
    out = np.empty(residual.shape, dtype=np.int32)
    ctx.check(ctx.lib.r1c_fwd_txfm_batch(ctx.h, tx_size, tx_type, bit_depth, _p(residual),
                                         _p(out), n))
    return out


def diff_forward_transform(ctx, plane_in, plane_pred, jobs, tx_size, tx_type, bit_depth):
    jobs = _arr(jobs, TX_JOB)
    w = ctx.lib.r1c_tx_width(tx_size)
    h = ctx.lib.r1c_tx_height(tx_size)
    out = np.empty((jobs.size, h, w), dtype=np.int32)
    ctx.check(ctx.lib.r1c_diff_fwd_txfm_batch(ctx.h, plane_in.h, plane_pred.h, _p(jobs),
                                              jobs.size, tx_size, tx_type, bit_depth, _p(out)))
    return out


def inverse_transform_add(ctx, rcoeffs, plane_dst, jobs, tx_size, tx_type, bit_depth, dequant=None):
    """inverse_transform_add (src/transform/mod.rs:568-580) over a batch: block i of `rcoeffs`
    ((N, coded) int32, coded = min(W,32)*min(H,32)) is inverse-transformed and added, clamped, to the
    W x H pixels at jobs[i] = (x, y) of the device plane `plane_dst`.  Blocks must not overlap.
    dequant = (dc_quant, ac_quant): `rcoeffs` holds QUANTIZED coefficients and dequantize is fused in."""
    jobs = _arr(jobs, ITX_JOB)
    rcoeffs = _arr(rcoeffs, np.int32)
    w, h = ctx.lib.r1c_tx_width(tx_size), ctx.lib.r1c_tx_height(tx_size)
    if w < 0:
        raise Rav1eCudaError(capi.R1C_ERR_ARG, "bad tx_size")
    assert rcoeffs.size == jobs.size * min(w, 32) * min(h, 32)
    if dequant is not None:
        ctx.check(ctx.lib.r1c_dequant_inv_txfm_add_batch(ctx.h, plane_dst.h, _p(jobs), jobs.size, tx_size, tx_type,
                                                         bit_depth, _p(rcoeffs), int(dequant[0]), int(dequant[1])))
        return
    ctx.check(ctx.lib.r1c_inv_txfm_add_batch(ctx.h, plane_dst.h, _p(jobs), jobs.size, tx_size, tx_type,
                                             bit_depth, _p(rcoeffs)))


def encode_tx_block_batch(ctx, plane_in, plane_rec, jobs, tx_size, tx_type, bit_depth, qc, scan, reconstruct=True,
                          want_tx_dist=True):
    """The residual chain of encode_tx_block (src/encoder.rs:1143-1224) for a batch of transform blocks as one kernel:
    diff of `plane_in` and the prediction held in `plane_rec` at jobs[i] -> forward transform -> quantize (qc: a
    QuantizationContext updated for tx_size; scan: the scan order) -> dequantize -> inverse transform added back onto
    `plane_rec` (when `reconstruct`) -> tx-domain distortion.  Returns (qcoeffs (N, min(W,32)*min(H,32)) int32 -- the
    coded area, av1_get_coded_tx_size -- and tx_dist (N,) uint64 or None).  All 19 TxSize."""
    jobs = _arr(jobs, TX_JOB)
    scan = _arr(scan, np.uint16)
    w, h = ctx.lib.r1c_tx_width(tx_size), ctx.lib.r1c_tx_height(tx_size)
    if w < 0:
        raise Rav1eCudaError(capi.R1C_ERR_ARG, "bad tx_size")
    assert scan.size == min(w, 32) * min(h, 32)
    q = np.empty((jobs.size, min(w, 32) * min(h, 32)), dtype=np.int32)
    d = np.empty(jobs.size, dtype=np.uint64) if want_tx_dist else None
    ctx.check(ctx.lib.r1c_encode_tx_block_batch(ctx.h, plane_in.h, plane_rec.h, _p(jobs), jobs.size, tx_size, tx_type,
                                                bit_depth, C.byref(qc.c), _p(scan), int(bool(reconstruct)), _p(q),
                                                _p(d) if d is not None else None))
    return q, d


# ----------------------------------------------------------------- motion search --
def get_mv_range(w_in_b, h_in_b, bo, blk_w, blk_h):
    out = (C.c_int64 * 4)()
    capi.load().r1c_mv_range(w_in_b, h_in_b, bo[0], bo[1], blk_w, blk_h, C.byref(out))
    return tuple(out)


def get_mv_rate(a, b, allow_high_precision_mv):
    return capi.load().r1c_mv_rate(capi.Mv(*a), capi.Mv(*b), int(allow_high_precision_mv))


def full_search_batch(ctx, plane_org, plane_ref, jobs, bit_depth, results=None):
    jobs = _arr(jobs, FS_JOB)
    if results is None:
        results = np.zeros(jobs.size, dtype=ME_RESULT)
        results["cost"] = np.iinfo(np.uint64).max
    results = _arr(results, ME_RESULT)
    ctx.check(ctx.lib.r1c_full_search_batch(ctx.h, plane_org.h, plane_ref.h, _p(jobs), jobs.size,
                                            bit_depth, _p(results)))
    return results


def full_search(ctx, x_lo, x_hi, y_lo, y_hi, bsize, p_org, p_ref, best_mv, lowest_cost, po, step,
                lambda_, pmv, allow_high_precision_mv, bit_depth):
    """Argument order of the reference's full_search (src/me.rs:943-949);
    returns the updated (best_mv, lowest_cost)."""
    j = np.zeros(1, dtype=FS_JOB)
    j["po_x"], j["po_y"], j["bsize"] = po[0], po[1], bsize
    j["x_lo"], j["x_hi"], j["y_lo"], j["y_hi"], j["step"] = x_lo, x_hi, y_lo, y_hi, step
    j["lambda_"] = lambda_
    j["pmv"][0, 0] = tuple(pmv[0])
    j["pmv"][0, 1] = tuple(pmv[1])
    j["allow_high_precision_mv"] = int(allow_high_precision_mv)
    j["has_init"] = 1
    r = np.zeros(1, dtype=ME_RESULT)
    r["mv"][0] = tuple(best_mv)
    r["cost"] = lowest_cost
    r = full_search_batch(ctx, p_org, p_ref, j, bit_depth, r)
    return (int(r["mv"]["row"][0]), int(r["mv"]["col"][0])), int(r["cost"][0])


def sb_result_offset(bsize):
    return capi.load().r1c_sb_result_offset(bsize)


def sb_result_count(bsize):
    return capi.load().r1c_sb_result_count(bsize)


def make_sb_params(w_in_b, h_in_b, range_x, range_y, step, lambda_, allow_hp=False):
    return capi.SbParams(w_in_b, h_in_b, range_x, range_y, step, lambda_, int(allow_hp))


def me_sb_batch(ctx, plane_org, plane_ref, params, jobs, bit_depth, results=None):
    jobs = _arr(jobs, SB_JOB)
    if results is None:
        results = np.zeros(jobs.size * SB_RESULTS, dtype=ME_RESULT)
        results["cost"] = np.iinfo(np.uint64).max
    ctx.check(ctx.lib.r1c_me_sb_batch(ctx.h, plane_org.h, plane_ref.h, C.byref(params),
                                      _p(jobs), jobs.size, bit_depth, _p(results)))
    return results.reshape(jobs.size, SB_RESULTS)


def subpel_cost_batch(ctx, plane_org, plane_ref, jobs, cand, bit_depth):
    jobs = _arr(jobs, SUBPEL_JOB)
    cand = _arr(cand, MV).reshape(jobs.size, -1)
    out = np.empty(cand.shape, dtype=np.uint64)
    ctx.check(ctx.lib.r1c_subpel_cost_batch(ctx.h, plane_org.h, plane_ref.h, _p(jobs), jobs.size,
                                            _p(cand), cand.shape[1], bit_depth, _p(out)))
    return out


def subpel_refine_batch(ctx, plane_org, plane_ref, jobs, results, bit_depth):
    jobs = _arr(jobs, SUBPEL_JOB)
    results = _arr(results, ME_RESULT).copy()
    ctx.check(ctx.lib.r1c_subpel_refine_batch(ctx.h, plane_org.h, plane_ref.h, _p(jobs),
                                              jobs.size, bit_depth, _p(results)))
    return results


def estimate_motion_ss4_batch(ctx, org_qres, ref_qres, w_in_b, h_in_b, jobs, me_range_scale, me_lambda, bit_depth,
                              allow_hp=False):
    """estimate_motion_ss4 (src/me.rs:1023-1075) for every BLOCK_JOB -> MV array (already x 4)"""
    jobs = _arr(jobs, BLOCK_JOB)
    out = np.zeros(jobs.size, dtype=MV)
    ctx.check(ctx.lib.r1c_estimate_motion_ss4_batch(ctx.h, org_qres.h, ref_qres.h, w_in_b, h_in_b, _p(jobs), jobs.size,
                                                    me_range_scale, float(me_lambda), int(allow_hp), bit_depth, _p(out)))
    return out


def estimate_motion_batch(ctx, org, ref, w_in_b, h_in_b, jobs, preds, valid, me_lambda, bit_depth, allow_hp=False, ss2=False):
    """ss2=True: estimate_motion_ss2 / FullSearch::me_ss2 on half-res planes, preds (n, 3); else estimate_motion /
    FullSearch::full_pixel_me on the full-res planes, preds (n, 7).  valid: same shape, non-zero = Some(mv)."""
    jobs = _arr(jobs, BLOCK_JOB)
    slots = 3 if ss2 else 7
    preds = _arr(preds, MV).reshape(jobs.size, slots)
    valid = _arr(valid, np.uint8).reshape(jobs.size, slots)
    out = np.zeros(jobs.size, dtype=MV)
    fn = ctx.lib.r1c_estimate_motion_ss2_batch if ss2 else ctx.lib.r1c_estimate_motion_batch
    ctx.check(fn(ctx.h, org.h, ref.h, w_in_b, h_in_b, _p(jobs), jobs.size, _p(preds), _p(valid), float(me_lambda),
                 int(allow_hp), bit_depth, _p(out)))
    return out


def save_block_motion(ctx, field, jobs, mvs):
    """save_block_motion (src/encoder.rs:1432-1444) for every BLOCK_JOB; field: MV array (rows, cols), updated in place"""
    jobs = _arr(jobs, BLOCK_JOB)
    mvs = _arr(mvs, MV)
    assert field.dtype == MV and field.flags.c_contiguous and field.ndim == 2 and mvs.size == jobs.size
    ctx.check(ctx.lib.r1c_save_block_motion_batch(ctx.h, _p(field), field.shape[1], field.shape[0], _p(jobs), _p(mvs), jobs.size))
    return field


def get_subset_predictors(ctx, field, prev_field, jobs, cmvs, n_cmvs):
    """get_subset_predictors (src/me.rs:82-174) -> (predictors (n, 17) MV, counts (n,))"""
    jobs = _arr(jobs, BLOCK_JOB)
    cmvs = _arr(cmvs, MV).reshape(jobs.size, 7)
    n_cmvs = _arr(n_cmvs, np.uint8)
    field = _arr(field, MV)
    pf = _arr(prev_field, MV) if prev_field is not None else None
    out = np.zeros((jobs.size, 17), dtype=MV)
    cnt = np.zeros(jobs.size, dtype=np.int32)
    ctx.check(ctx.lib.r1c_subset_predictors_batch(ctx.h, _p(field), _p(pf) if pf is not None else None, field.shape[1],
                                                  field.shape[0], _p(jobs), jobs.size, _p(cmvs), _p(n_cmvs), _p(out), _p(cnt)))
    return out, cnt


def diamond_search_batch(ctx, plane_org, plane_ref, jobs, predictors, bit_depth, subpixel=False):
    """diamond_me_search + get_best_predictor (src/me.rs:655-785) for every SUBPEL_JOB;
    predictors: MV array (n, n_pred).  Returns ME_RESULT (centre MV, cost) per job."""
    jobs = _arr(jobs, SUBPEL_JOB)
    predictors = _arr(predictors, MV).reshape(jobs.size, -1)
    out = np.zeros(jobs.size, dtype=ME_RESULT)
    ctx.check(ctx.lib.r1c_diamond_search_batch(ctx.h, plane_org.h, plane_ref.h, _p(jobs), jobs.size, int(subpixel),
                                               _p(predictors), predictors.shape[1], bit_depth, _p(out)))
    return out


# ----------------------------------------------------- whole-frame motion search --
ALL_SIZES_TO_64 = sum(1 << b for b in range(22) if b not in (13, 14, 15))


def make_frame_me_params(width, height, range_x, range_y, lambda_, bsize_mask=ALL_SIZES_TO_64,
                         refine=1, use_satd=1, allow_hp=False, sb_rows=(0, 0), sb_cols=(0, 0)):
    """sb_rows / sb_cols: the tile (superblock rows [lo, hi) x columns [lo, hi)) a sharded call computes; (0, 0) = all"""
    return capi.FrameMeParams(width, height, range_x, range_y, lambda_, bsize_mask, int(refine),
                              int(use_satd), int(allow_hp), sb_rows[0], sb_rows[1], sb_cols[0], sb_cols[1])


def frame_me_window(ctx, pairs, params, bit_depth):
    """r1c_frame_me_window: a lookahead window of frame pairs in one call, alternating over two streams.
    pairs: list of (plane_org, plane_ref) or (plane_org, plane_ref, org_host, ref_host) with numpy arrays of the whole
    padded planes to upload first.  Returns the list of ME_RESULT arrays (one per pair)."""
    n = frame_me_count(params)
    if n < 0:
        raise Rav1eCudaError(capi.R1C_ERR_ARG, "bad frame_me parameters")
    arr = (capi.FramePair * max(len(pairs), 1))()
    outs, keep = [], []
    for i, pr in enumerate(pairs):
        res = np.zeros(n, dtype=ME_RESULT)
        outs.append(res)
        arr[i].org, arr[i].ref, arr[i].results = pr[0].h, pr[1].h, res.ctypes.data
        if len(pr) > 2:
            ho, hr = np.ascontiguousarray(pr[2]), np.ascontiguousarray(pr[3])
            keep += [ho, hr]
            arr[i].org_host, arr[i].ref_host = ho.ctypes.data, hr.ctypes.data
    ctx.check(ctx.lib.r1c_frame_me_window(ctx.h, arr, len(pairs), C.byref(params), bit_depth))
    return outs


def frame_me_count(params):
    return capi.load().r1c_frame_me_count(C.byref(params))


def frame_me_layout(params, bsize):
    """-> (offset, cols, rows) of one BlockSize inside the frame_me result array"""
    cols, rows = C.c_int32(), C.c_int32()
    off = capi.load().r1c_frame_me_offset(C.byref(params), bsize, C.byref(cols), C.byref(rows))
    return off, cols.value, rows.value


def frame_me(ctx, plane_org, plane_ref, params, bit_depth, results=None):
    """MotionEstimation::motion_estimation (FullSearch, src/me.rs:193-271) for every block of
    every size of the frame; `results` (ME_RESULT array, host) is in/out for sharded calls."""
    n = frame_me_count(params)
    if n < 0:
        raise Rav1eCudaError(capi.R1C_ERR_ARG, "bad frame_me parameters")
    if results is None:
        results = np.zeros(n, dtype=ME_RESULT)
        results["cost"] = np.iinfo(np.uint64).max
    assert results.dtype == ME_RESULT and results.size == n and results.flags.c_contiguous
    ctx.check(ctx.lib.r1c_frame_me(ctx.h, plane_org.h, plane_ref.h, C.byref(params), bit_depth,
                                   _p(results)))
    return results
