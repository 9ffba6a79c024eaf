"""ctypes declarations of include/rav1e_cuda.h (librav1e_cuda.so).

This is the thin host-side binding used by the Python mirror of the
reference interface (rav1e_b200/api.py), by the parity tests and by bench.py.
It loads the in-tree shared library and fails loudly when it is missing:
there is no CPU fallback anywhere in this package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librav1e_cuda.so")

R1C_OK = 0
R1C_ERR_ARG = -1
R1C_ERR_CUDA = -2
R1C_ERR_BOUNDS = -3
R1C_ERR_UNIMPLEMENTED = -4
R1C_ERR_NOMEM = -5
R1C_SB_RESULTS = 849
R1C_SB_NOALLOC_SCRATCH_BYTES = 8 << 20

BLOCK_SIZES = [
    "BLOCK_4X4", "BLOCK_4X8", "BLOCK_8X4", "BLOCK_8X8", "BLOCK_8X16", "BLOCK_16X8",
    "BLOCK_16X16", "BLOCK_16X32", "BLOCK_32X16", "BLOCK_32X32", "BLOCK_32X64", "BLOCK_64X32",
    "BLOCK_64X64", "BLOCK_64X128", "BLOCK_128X64", "BLOCK_128X128", "BLOCK_4X16", "BLOCK_16X4",
    "BLOCK_8X32", "BLOCK_32X8", "BLOCK_16X64", "BLOCK_64X16"]
TX_SIZES = [
    "TX_4X4", "TX_8X8", "TX_16X16", "TX_32X32", "TX_64X64", "TX_4X8", "TX_8X4", "TX_8X16",
    "TX_16X8", "TX_16X32", "TX_32X16", "TX_32X64", "TX_64X32", "TX_4X16", "TX_16X4", "TX_8X32",
    "TX_32X8", "TX_16X64", "TX_64X16"]
TX_TYPES = [
    "DCT_DCT", "ADST_DCT", "DCT_ADST", "ADST_ADST", "FLIPADST_DCT", "DCT_FLIPADST",
    "FLIPADST_FLIPADST", "ADST_FLIPADST", "FLIPADST_ADST", "IDTX", "V_DCT", "H_DCT", "V_ADST",
    "H_ADST", "V_FLIPADST", "H_FLIPADST"]
FILTER_MODES = ["REGULAR", "SMOOTH", "SHARP", "BILINEAR"]


class PlaneCfg(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "stride", "alloc_height", "width", "height", "xdec", "ydec", "xpad", "ypad",
        "xorigin", "yorigin")]


class DistJob(C.Structure):
    _fields_ = [("org_x", C.c_int32), ("org_y", C.c_int32), ("ref_x", C.c_int32),
                ("ref_y", C.c_int32), ("bsize", C.c_int32)]


class McJob(C.Structure):
    _fields_ = [("dst_x", C.c_int32), ("dst_y", C.c_int32), ("src_x", C.c_int32),
                ("src_y", C.c_int32), ("w", C.c_int16), ("h", C.c_int16),
                ("col_frac", C.c_int8), ("row_frac", C.c_int8), ("mode_x", C.c_int8),
                ("mode_y", C.c_int8), ("tmp_off", C.c_int32)]


class AvgJob(C.Structure):
    _fields_ = [("dst_x", C.c_int32), ("dst_y", C.c_int32), ("w", C.c_int16), ("h", C.c_int16),
                ("tmp_off", C.c_int32)]


class TxJob(C.Structure):
    _fields_ = [("in_x", C.c_int32), ("in_y", C.c_int32), ("pred_x", C.c_int32),
                ("pred_y", C.c_int32)]


class ItxJob(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32)]


class BlockJob(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("bo_x", "bo_y", "bsize")]


class RdoJob(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("src_x", "src_y", "rec_x", "rec_y", "w", "h")]


class BiasParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("w_in_imp_b", "n_imp", "w_in_b", "h_in_b")]


class QuantCtx(C.Structure):
    """r1c_quant_ctx = QuantizationContext, src/quantize.rs:111-121"""
    _fields_ = [("log_tx_scale", C.c_uint32), ("dc_quant", C.c_uint32), ("ac_quant", C.c_uint32),
                ("dc_offset", C.c_int32), ("ac_offset_eob", C.c_int32), ("ac_offset0", C.c_int32),
                ("ac_offset1", C.c_int32), ("dc_mul_add", C.c_uint32 * 3), ("ac_mul_add", C.c_uint32 * 3)]


class Mv(C.Structure):
    _fields_ = [("row", C.c_int16), ("col", C.c_int16)]


class MeResult(C.Structure):
    _fields_ = [("mv", Mv), ("reserved", C.c_uint32), ("cost", C.c_uint64)]


class FsJob(C.Structure):
    _fields_ = [("po_x", C.c_int32), ("po_y", C.c_int32), ("bsize", C.c_int32),
                ("x_lo", C.c_int32), ("x_hi", C.c_int32), ("y_lo", C.c_int32),
                ("y_hi", C.c_int32), ("step", C.c_int32), ("lambda_", C.c_uint32),
                ("pmv", Mv * 2), ("allow_high_precision_mv", C.c_int32),
                ("has_init", C.c_int32)]


class SbJob(C.Structure):
    _fields_ = [("po_x", C.c_int32), ("po_y", C.c_int32), ("bsize_mask", C.c_uint32),
                ("cmv", Mv), ("pmv", Mv * 2)]


class SbParams(C.Structure):
    _fields_ = [("w_in_b", C.c_int32), ("h_in_b", C.c_int32), ("range_x", C.c_int32),
                ("range_y", C.c_int32), ("step", C.c_int32), ("lambda_", C.c_uint32),
                ("allow_high_precision_mv", C.c_int32)]


class SubpelJob(C.Structure):
    _fields_ = [("po_x", C.c_int32), ("po_y", C.c_int32), ("bsize", C.c_int32),
                ("lambda_", C.c_uint32), ("pmv", Mv * 2), ("mvx_min", C.c_int32),
                ("mvx_max", C.c_int32), ("mvy_min", C.c_int32), ("mvy_max", C.c_int32),
                ("use_satd", C.c_int32), ("allow_high_precision_mv", C.c_int32)]


class FramePair(C.Structure):
    _fields_ = [("org", C.c_void_p), ("ref", C.c_void_p), ("org_host", C.c_void_p), ("ref_host", C.c_void_p),
                ("results", C.c_void_p)]


class FrameMeParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("range_x", C.c_int32),
                ("range_y", C.c_int32), ("lambda_", C.c_uint32), ("bsize_mask", C.c_uint32),
                ("refine", C.c_int32), ("use_satd", C.c_int32),
                ("allow_high_precision_mv", C.c_int32), ("sb_row_lo", C.c_int32),
                ("sb_row_hi", C.c_int32), ("sb_col_lo", C.c_int32), ("sb_col_hi", C.c_int32)]


# every symbol include/rav1e_cuda.h declares: name -> (restype, argtypes)
_vp, _i, _i64, _u32, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64
_P = C.POINTER
SIGNATURES = {
    "r1c_block_width": (_i, [_i]), "r1c_block_height": (_i, [_i]),
    "r1c_tx_width": (_i, [_i]), "r1c_tx_height": (_i, [_i]),
    "r1c_fwd_txfm_valid": (_i, [_i, _i]),
    "r1c_create": (_vp, [_i]), "r1c_destroy": (None, [_vp]), "r1c_sync": (_i, [_vp]),
    "r1c_last_error": (C.c_char_p, [_vp]), "r1c_stream": (_vp, [_vp]),
    "r1c_device": (_i, [_vp]), "r1c_launch_count": (_i64, [_vp]),
    "r1c_plane_cfg_new": (_i, [_P(PlaneCfg), _i64, _i64, _i64, _i64, _i64, _i64, _i]),
    "r1c_plane_create": (_i, [_vp, _P(PlaneCfg), _i, _P(_vp)]),
    "r1c_plane_wrap": (_i, [_vp, _vp, _P(PlaneCfg), _i, _P(_vp)]),
    "r1c_plane_destroy": (None, [_vp]),
    "r1c_plane_upload": (_i, [_vp, _vp, _vp]), "r1c_plane_download": (_i, [_vp, _vp, _vp]),
    "r1c_plane_device_ptr": (_vp, [_vp]), "r1c_plane_config": (_P(PlaneCfg), [_vp]),
    "r1c_plane_pad": (_i, [_vp, _vp, _i64, _i64]),
    "r1c_plane_downsample": (_i, [_vp, _vp, _vp]),
    "r1c_sad": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _i, _i, _P(_u32)]),
    "r1c_satd": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _i, _i, _P(_u32)]),
    "r1c_sad_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_satd_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_satd_i16_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_lookahead_intra_costs": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "r1c_lookahead_intra_costs_dev": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "r1c_dist_batch_dev": (_i, [_vp, _i, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_mv_params": (None, [_P(PlaneCfg), _i64, _i64, _i, _i, _P(_i), _P(_i), _P(_i64), _P(_i64)]),
    "r1c_put_8tap_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i]),
    "r1c_prep_8tap_batch": (_i, [_vp, _vp, _i64, _vp, _vp, _i64, _i]),
    "r1c_mc_avg_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _i64, _i]),
    "r1c_put_8tap_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i]),
    "r1c_prep_8tap_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i]),
    "r1c_mc_avg_batch_dev": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i]),
    "r1c_forward_transform": (_i, [_vp, _vp, _vp, _i, _i, _i]),
    "r1c_fwd_txfm_batch": (_i, [_vp, _i, _i, _i, _vp, _vp, _i64]),
    "r1c_fwd_txfm_batch_dev": (_i, [_vp, _i, _i, _i, _vp, _vp, _i64]),
    "r1c_sse_wxh_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp, _vp, _vp]),
    "r1c_cdef_dist_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp, _vp, _vp]),
    "r1c_rdo_dist_batch_dev": (_i, [_vp, _i, _vp, _vp, _vp, _i64, _i, _vp, _vp, _vp]),
    "r1c_tx_dist_batch": (_i, [_vp, _vp, _vp, _i, _i64, _vp]),
    "r1c_tx_dist_batch_dev": (_i, [_vp, _vp, _vp, _i, _i64, _vp]),
    "r1c_importance_propagate": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    "r1c_importance_propagate_dev": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    "r1c_importance_finalize": (_i, [_vp, _vp, _vp, _i64]),
    "r1c_importance_finalize_dev": (_i, [_vp, _vp, _vp, _i64]),
    "r1c_log_tx_scale": (_i, [_i]),
    "r1c_quant_ctx_update": (_i, [_P(QuantCtx), _i, _i, _u32, _u32]),
    "r1c_quantize_batch": (_i, [_vp, _P(QuantCtx), _i, _vp, _vp, _vp, _i64]),
    "r1c_quantize_batch_dev": (_i, [_vp, _P(QuantCtx), _i, _vp, _vp, _vp, _vp, _vp, _i64]),
    "r1c_dequantize_batch": (_i, [_vp, _i, _u32, _u32, _vp, _vp, _i64]),
    "r1c_dequantize_batch_dev": (_i, [_vp, _i, _u32, _u32, _vp, _vp, _i64]),
    "r1c_diff_fwd_txfm_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "r1c_diff_fwd_txfm_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "r1c_inv_txfm_valid": (_i, [_i, _i]),
    "r1c_inv_txfm_add_batch": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "r1c_inv_txfm_add_batch_dev": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "r1c_dequant_inv_txfm_add_batch": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _u32, _u32]),
    "r1c_dequant_inv_txfm_add_batch_dev": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _u32, _u32]),
    "r1c_encode_tx_block_valid": (_i, [_i, _i]),
    "r1c_encode_tx_block_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _P(QuantCtx), _vp, _i, _vp, _vp]),
    "r1c_encode_tx_block_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _P(QuantCtx), _vp, _i, _vp, _vp]),
    "r1c_mv_range": (None, [_i64, _i64, _i64, _i64, _i, _i, _P(_i64 * 4)]),
    "r1c_mv_rate": (_u32, [Mv, Mv, _i]),
    "r1c_full_search_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_full_search_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_sb_result_offset": (_i, [_i]), "r1c_sb_result_count": (_i, [_i]),
    "r1c_me_sb_batch": (_i, [_vp, _vp, _vp, _P(SbParams), _vp, _i64, _i, _vp]),
    "r1c_me_sb_batch_dev": (_i, [_vp, _vp, _vp, _P(SbParams), _vp, _i64, _i, _vp]),
    "r1c_me_sb_batch_dev_noalloc": (_i, [_vp, _vp, _vp, _P(SbParams), _vp, _i64, _i, _vp]),
    "r1c_subpel_cost_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _i, _i, _vp]),
    "r1c_subpel_refine_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_subpel_refine_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "r1c_estimate_motion_ss4_batch": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i64, _i, C.c_double, _i, _i, _vp]),
    "r1c_estimate_motion_ss2_batch": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i64, _vp, _vp, C.c_double, _i, _i, _vp]),
    "r1c_estimate_motion_batch": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i64, _vp, _vp, C.c_double, _i, _i, _vp]),
    "r1c_save_block_motion_batch": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i64]),
    "r1c_save_block_motion_batch_dev": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i64]),
    "r1c_subset_predictors_batch": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i64, _vp, _vp, _vp, _vp]),
    "r1c_subset_predictors_batch_dev": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i64, _vp, _vp, _vp, _vp]),
    "r1c_diamond_search_batch": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp, _i, _i, _vp]),
    "r1c_diamond_search_batch_dev": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _vp, _i, _i, _vp]),
    "r1c_frame_me_count": (_i64, [_P(FrameMeParams)]),
    "r1c_frame_me_offset": (_i64, [_P(FrameMeParams), _i, _P(C.c_int32), _P(C.c_int32)]),
    "r1c_frame_me": (_i, [_vp, _vp, _vp, _P(FrameMeParams), _i, _vp]),
    "r1c_frame_me_dev": (_i, [_vp, _vp, _vp, _P(FrameMeParams), _i, _vp]),
    "r1c_frame_me_window": (_i, [_vp, _vp, _i64, _P(FrameMeParams), _i]),
    "r1c_frame_me_window_dev": (_i, [_vp, _vp, _i64, _P(FrameMeParams), _i]),
    "r1c_frame_me_shard_count": (_i64, [_vp, _P(FrameMeParams)]),
    "r1c_frame_me_compact_dev": (_i, [_vp, _P(FrameMeParams), _vp, _vp]),
    "r1c_plane_upload_async": (_i, [_vp, _vp, _vp]),
    "r1c_plane_upload_rows_async": (_i, [_vp, _vp, _vp, _i64, _i64]),
    "r1c_plane_upload_rect_async": (_i, [_vp, _vp, _vp, _i64, _i64, _i64, _i64]),
    "r1c_int_peak": (_i, [_vp, _i, _P(C.c_double)]),
    "r1c_profile": (_i, [_vp, _i]),
    "r1c_profile_read": (_i, [_vp, _i, _P(C.c_double), _P(_i64)]),
}

_lib = None


class Rav1eCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rav1e_cuda error %d: %s" % (code, msg))
        self.code = code


def load():
    """dlopen the in-tree librav1e_cuda.so and bind every header symbol."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). rav1e_b200 has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
