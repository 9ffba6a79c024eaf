// rav1e_cuda_sys.rs -- GENERATED from include/rav1e_cuda.h by tools/gen_rust_ffi.py; do not edit.
// The `extern "C"` surface of librav1e_cuda.so as the `rav1e-cuda-sys` crate of INTEGRATION.md section 1.
#![allow(non_camel_case_types, non_upper_case_globals, dead_code)]
use std::os::raw::{c_char, c_int, c_void};

pub const R1C_OK: c_int = 0;
pub const R1C_ERR_ARG: c_int = -1;
pub const R1C_ERR_CUDA: c_int = -2;
pub const R1C_ERR_BOUNDS: c_int = -3;
pub const R1C_ERR_UNIMPLEMENTED: c_int = -4;
pub const R1C_ERR_NOMEM: c_int = -5;
pub const R1C_PROF_ME_SB: c_int = 0;
pub const R1C_PROF_SUBPEL_MC: c_int = 1;
pub const R1C_PROF_SUBPEL_COST: c_int = 2;
pub const R1C_PROF_MC: c_int = 3;
pub const R1C_PROF_TX: c_int = 4;
pub const R1C_PROF_DIST: c_int = 5;
pub const R1C_PROF_QUANT: c_int = 6;
pub const R1C_PROF_ITX: c_int = 7;
pub const R1C_PROF_SLOTS: c_int = 8;
pub const R1C_PEAK_IADD: c_int = 0;
pub const R1C_PEAK_VABSDIFF4: c_int = 1;
pub const R1C_PEAK_DP4A: c_int = 2;
pub const R1C_PEAK_DP2A: c_int = 3;
pub const R1C_PEAK_IMAD: c_int = 4;
pub const R1C_PEAK_PRMT: c_int = 5;
pub const R1C_DIST_SAD: c_int = 0;
pub const R1C_DIST_SATD: c_int = 1;
pub const R1C_DIST_SATD_I16: c_int = 2;
pub const R1C_SB_RESULTS: c_int = 849;
pub const R1C_SB_NOALLOC_SCRATCH_BYTES: u32 = 8388608;

pub type r1c_block_size = c_int;
pub const R1C_BLOCK_4X4: r1c_block_size = 0;
pub const R1C_BLOCK_4X8: r1c_block_size = 1;
pub const R1C_BLOCK_8X4: r1c_block_size = 2;
pub const R1C_BLOCK_8X8: r1c_block_size = 3;
pub const R1C_BLOCK_8X16: r1c_block_size = 4;
pub const R1C_BLOCK_16X8: r1c_block_size = 5;
pub const R1C_BLOCK_16X16: r1c_block_size = 6;
pub const R1C_BLOCK_16X32: r1c_block_size = 7;
pub const R1C_BLOCK_32X16: r1c_block_size = 8;
pub const R1C_BLOCK_32X32: r1c_block_size = 9;
pub const R1C_BLOCK_32X64: r1c_block_size = 10;
pub const R1C_BLOCK_64X32: r1c_block_size = 11;
pub const R1C_BLOCK_64X64: r1c_block_size = 12;
pub const R1C_BLOCK_64X128: r1c_block_size = 13;
pub const R1C_BLOCK_128X64: r1c_block_size = 14;
pub const R1C_BLOCK_128X128: r1c_block_size = 15;
pub const R1C_BLOCK_4X16: r1c_block_size = 16;
pub const R1C_BLOCK_16X4: r1c_block_size = 17;
pub const R1C_BLOCK_8X32: r1c_block_size = 18;
pub const R1C_BLOCK_32X8: r1c_block_size = 19;
pub const R1C_BLOCK_16X64: r1c_block_size = 20;
pub const R1C_BLOCK_64X16: r1c_block_size = 21;
pub const R1C_BLOCK_SIZES_ALL: r1c_block_size = 22;

pub type r1c_filter_mode = c_int;
pub const R1C_REGULAR: r1c_filter_mode = 0;
pub const R1C_SMOOTH: r1c_filter_mode = 1;
pub const R1C_SHARP: r1c_filter_mode = 2;
pub const R1C_BILINEAR: r1c_filter_mode = 3;

pub type r1c_tx_size = c_int;
pub const R1C_TX_4X4: r1c_tx_size = 0;
pub const R1C_TX_8X8: r1c_tx_size = 1;
pub const R1C_TX_16X16: r1c_tx_size = 2;
pub const R1C_TX_32X32: r1c_tx_size = 3;
pub const R1C_TX_64X64: r1c_tx_size = 4;
pub const R1C_TX_4X8: r1c_tx_size = 5;
pub const R1C_TX_8X4: r1c_tx_size = 6;
pub const R1C_TX_8X16: r1c_tx_size = 7;
pub const R1C_TX_16X8: r1c_tx_size = 8;
pub const R1C_TX_16X32: r1c_tx_size = 9;
pub const R1C_TX_32X16: r1c_tx_size = 10;
pub const R1C_TX_32X64: r1c_tx_size = 11;
pub const R1C_TX_64X32: r1c_tx_size = 12;
pub const R1C_TX_4X16: r1c_tx_size = 13;
pub const R1C_TX_16X4: r1c_tx_size = 14;
pub const R1C_TX_8X32: r1c_tx_size = 15;
pub const R1C_TX_32X8: r1c_tx_size = 16;
pub const R1C_TX_16X64: r1c_tx_size = 17;
pub const R1C_TX_64X16: r1c_tx_size = 18;
pub const R1C_TX_SIZES_ALL: r1c_tx_size = 19;

pub type r1c_tx_type = c_int;
pub const R1C_DCT_DCT: r1c_tx_type = 0;
pub const R1C_ADST_DCT: r1c_tx_type = 1;
pub const R1C_DCT_ADST: r1c_tx_type = 2;
pub const R1C_ADST_ADST: r1c_tx_type = 3;
pub const R1C_FLIPADST_DCT: r1c_tx_type = 4;
pub const R1C_DCT_FLIPADST: r1c_tx_type = 5;
pub const R1C_FLIPADST_FLIPADST: r1c_tx_type = 6;
pub const R1C_ADST_FLIPADST: r1c_tx_type = 7;
pub const R1C_FLIPADST_ADST: r1c_tx_type = 8;
pub const R1C_IDTX: r1c_tx_type = 9;
pub const R1C_V_DCT: r1c_tx_type = 10;
pub const R1C_H_DCT: r1c_tx_type = 11;
pub const R1C_V_ADST: r1c_tx_type = 12;
pub const R1C_H_ADST: r1c_tx_type = 13;
pub const R1C_V_FLIPADST: r1c_tx_type = 14;
pub const R1C_H_FLIPADST: r1c_tx_type = 15;
pub const R1C_TX_TYPES: r1c_tx_type = 16;

#[repr(C)] pub struct r1c_ctx { _private: [u8; 0] }
#[repr(C)] pub struct r1c_plane { _private: [u8; 0] }

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_plane_cfg {
  pub stride: i64,
  pub alloc_height: i64,
  pub width: i64,
  pub height: i64,
  pub xdec: i64,
  pub ydec: i64,
  pub xpad: i64,
  pub ypad: i64,
  pub xorigin: i64,
  pub yorigin: i64,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_dist_job {
  pub org_x: i32,
  pub org_y: i32,
  pub ref_x: i32,
  pub ref_y: i32,
  pub bsize: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_mc_job {
  pub dst_x: i32,
  pub dst_y: i32,
  pub src_x: i32,
  pub src_y: i32,
  pub w: i16,
  pub h: i16,
  pub col_frac: i8,
  pub row_frac: i8,
  pub mode_x: i8,
  pub mode_y: i8,
  pub tmp_off: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_avg_job {
  pub dst_x: i32,
  pub dst_y: i32,
  pub w: i16,
  pub h: i16,
  pub tmp_off: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_rdo_job {
  pub src_x: i32,
  pub src_y: i32,
  pub rec_x: i32,
  pub rec_y: i32,
  pub w: i32,
  pub h: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_bias_params {
  pub w_in_imp_b: i32,
  pub n_imp: i32,
  pub w_in_b: i32,
  pub h_in_b: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_tx_job {
  pub in_x: i32,
  pub in_y: i32,
  pub pred_x: i32,
  pub pred_y: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_itx_job {
  pub x: i32,
  pub y: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_quant_ctx {
  pub log_tx_scale: u32,
  pub dc_quant: u32,
  pub ac_quant: u32,
  pub dc_offset: i32,
  pub ac_offset_eob: i32,
  pub ac_offset0: i32,
  pub ac_offset1: i32,
  pub dc_mul_add: [u32; 3],
  pub ac_mul_add: [u32; 3],
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_mv {
  pub row: i16,
  pub col: i16,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_me_result {
  pub mv: r1c_mv,
  pub reserved: u32,
  pub cost: u64,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_fs_job {
  pub po_x: i32,
  pub po_y: i32,
  pub bsize: i32,
  pub x_lo: i32,
  pub x_hi: i32,
  pub y_lo: i32,
  pub y_hi: i32,
  pub step: i32,
  pub lambda: u32,
  pub pmv: [r1c_mv; 2],
  pub allow_high_precision_mv: i32,
  pub has_init: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_sb_job {
  pub po_x: i32,
  pub po_y: i32,
  pub bsize_mask: u32,
  pub cmv: r1c_mv,
  pub pmv: [r1c_mv; 2],
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_sb_params {
  pub w_in_b: i32,
  pub h_in_b: i32,
  pub range_x: i32,
  pub range_y: i32,
  pub step: i32,
  pub lambda: u32,
  pub allow_high_precision_mv: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_block_job {
  pub bo_x: i32,
  pub bo_y: i32,
  pub bsize: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_subpel_job {
  pub po_x: i32,
  pub po_y: i32,
  pub bsize: i32,
  pub lambda: u32,
  pub pmv: [r1c_mv; 2],
  pub mvx_min: i32,
  pub mvx_max: i32,
  pub mvy_min: i32,
  pub mvy_max: i32,
  pub use_satd: i32,
  pub allow_high_precision_mv: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_frame_me_params {
  pub width: i32,
  pub height: i32,
  pub range_x: i32,
  pub range_y: i32,
  pub lambda: u32,
  pub bsize_mask: u32,
  pub refine: i32,
  pub use_satd: i32,
  pub allow_high_precision_mv: i32,
  pub sb_row_lo: i32,
  pub sb_row_hi: i32,
  pub sb_col_lo: i32,
  pub sb_col_hi: i32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct r1c_frame_pair {
  pub org: *const r1c_plane,
  pub r#ref: *const r1c_plane,
  pub org_host: *const c_void,
  pub ref_host: *const c_void,
  pub results: *mut c_void,
}

#[link(name = "rav1e_cuda")]
extern "C" {
  pub fn r1c_block_width(bsize: c_int) -> c_int;
  pub fn r1c_block_height(bsize: c_int) -> c_int;
  pub fn r1c_tx_width(tx_size: c_int) -> c_int;
  pub fn r1c_tx_height(tx_size: c_int) -> c_int;
  pub fn r1c_fwd_txfm_valid(tx_size: c_int, tx_type: c_int) -> c_int;
  pub fn r1c_create(device: c_int) -> *mut r1c_ctx;
  pub fn r1c_destroy(ctx: *mut r1c_ctx);
  pub fn r1c_sync(ctx: *mut r1c_ctx) -> c_int;
  pub fn r1c_last_error(ctx: *const r1c_ctx) -> *const c_char;
  pub fn r1c_stream(ctx: *mut r1c_ctx) -> *mut c_void;
  pub fn r1c_device(ctx: *const r1c_ctx) -> c_int;
  pub fn r1c_launch_count(ctx: *const r1c_ctx) -> i64;
  pub fn r1c_int_peak(ctx: *mut r1c_ctx, which: c_int, giga_lane_ops_per_s: *mut f64) -> c_int;
  pub fn r1c_profile(ctx: *mut r1c_ctx, enable: c_int) -> c_int;
  pub fn r1c_profile_read(ctx: *mut r1c_ctx, slot: c_int, ms: *mut f64, launches: *mut i64) -> c_int;
  pub fn r1c_plane_cfg_new(cfg: *mut r1c_plane_cfg, width: i64, height: i64, xdec: i64, ydec: i64, xpad: i64, ypad: i64, px_bytes: c_int) -> c_int;
  pub fn r1c_plane_create(ctx: *mut r1c_ctx, cfg: *const r1c_plane_cfg, px_bytes: c_int, out: *mut *mut r1c_plane) -> c_int;
  pub fn r1c_plane_wrap(ctx: *mut r1c_ctx, device_data: *mut c_void, cfg: *const r1c_plane_cfg, px_bytes: c_int, out: *mut *mut r1c_plane) -> c_int;
  pub fn r1c_plane_destroy(plane: *mut r1c_plane);
  pub fn r1c_plane_upload(ctx: *mut r1c_ctx, plane: *mut r1c_plane, host_data: *const c_void) -> c_int;
  pub fn r1c_plane_download(ctx: *mut r1c_ctx, plane: *const r1c_plane, host_data: *mut c_void) -> c_int;
  pub fn r1c_plane_device_ptr(plane: *const r1c_plane) -> *mut c_void;
  pub fn r1c_plane_config(plane: *const r1c_plane) -> *const r1c_plane_cfg;
  pub fn r1c_plane_pad(ctx: *mut r1c_ctx, plane: *mut r1c_plane, w: i64, h: i64) -> c_int;
  pub fn r1c_plane_downsample(ctx: *mut r1c_ctx, dst: *mut r1c_plane, src: *const r1c_plane) -> c_int;
  pub fn r1c_sad(ctx: *mut r1c_ctx, org: *const r1c_plane, org_x: c_int, org_y: c_int, r#ref: *const r1c_plane, ref_x: c_int, ref_y: c_int, bsize: c_int, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_satd(ctx: *mut r1c_ctx, org: *const r1c_plane, org_x: c_int, org_y: c_int, r#ref: *const r1c_plane, ref_x: c_int, ref_y: c_int, bsize: c_int, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_sad_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_dist_job, n: i64, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_satd_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_dist_job, n: i64, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_dist_batch_dev(ctx: *mut r1c_ctx, use_satd: c_int, org: *const r1c_plane, r#ref: *const r1c_plane, d_jobs: *const r1c_dist_job, n: i64, bit_depth: c_int, d_out: *mut u32) -> c_int;
  pub fn r1c_satd_i16_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_dist_job, n: i64, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_lookahead_intra_costs(ctx: *mut r1c_ctx, input: *const r1c_plane, w_in_imp_b: c_int, h_in_imp_b: c_int, bit_depth: c_int, out: *mut u32) -> c_int;
  pub fn r1c_lookahead_intra_costs_dev(ctx: *mut r1c_ctx, input: *const r1c_plane, w_in_imp_b: c_int, h_in_imp_b: c_int, bit_depth: c_int, d_out: *mut u32) -> c_int;
  pub fn r1c_mv_params(rec_cfg: *const r1c_plane_cfg, po_x: i64, po_y: i64, mv_row: c_int, mv_col: c_int, row_frac: *mut c_int, col_frac: *mut c_int, src_x: *mut i64, src_y: *mut i64);
  pub fn r1c_put_8tap_batch(ctx: *mut r1c_ctx, dst: *mut r1c_plane, src: *const r1c_plane, jobs: *const r1c_mc_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_prep_8tap_batch(ctx: *mut r1c_ctx, tmp: *mut i16, tmp_len: i64, src: *const r1c_plane, jobs: *const r1c_mc_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_mc_avg_batch(ctx: *mut r1c_ctx, dst: *mut r1c_plane, tmp1: *const i16, tmp2: *const i16, tmp_len: i64, jobs: *const r1c_avg_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_put_8tap_batch_dev(ctx: *mut r1c_ctx, dst: *mut r1c_plane, src: *const r1c_plane, d_jobs: *const r1c_mc_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_prep_8tap_batch_dev(ctx: *mut r1c_ctx, d_tmp: *mut i16, src: *const r1c_plane, d_jobs: *const r1c_mc_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_mc_avg_batch_dev(ctx: *mut r1c_ctx, dst: *mut r1c_plane, d_tmp1: *const i16, d_tmp2: *const i16, d_jobs: *const r1c_avg_job, n: i64, bit_depth: c_int) -> c_int;
  pub fn r1c_sse_wxh_batch(ctx: *mut r1c_ctx, src1: *const r1c_plane, src2: *const r1c_plane, jobs: *const r1c_rdo_job, n: i64, bit_depth: c_int, block_importances: *const f32, bp: *const r1c_bias_params, out: *mut u64) -> c_int;
  pub fn r1c_cdef_dist_batch(ctx: *mut r1c_ctx, src1: *const r1c_plane, src2: *const r1c_plane, jobs: *const r1c_rdo_job, n: i64, bit_depth: c_int, block_importances: *const f32, bp: *const r1c_bias_params, out: *mut u64) -> c_int;
  pub fn r1c_rdo_dist_batch_dev(ctx: *mut r1c_ctx, kind: c_int, src1: *const r1c_plane, src2: *const r1c_plane, d_jobs: *const r1c_rdo_job, n: i64, bit_depth: c_int, d_block_importances: *const f32, bp: *const r1c_bias_params, d_out: *mut u64) -> c_int;
  pub fn r1c_tx_dist_batch(ctx: *mut r1c_ctx, coeffs: *const i32, rcoeffs: *const i32, tx_size: c_int, n: i64, out: *mut u64) -> c_int;
  pub fn r1c_tx_dist_batch_dev(ctx: *mut r1c_ctx, d_coeffs: *const i32, d_rcoeffs: *const i32, tx_size: c_int, n: i64, d_out: *mut u64) -> c_int;
  pub fn r1c_forward_transform(ctx: *mut r1c_ctx, residual: *const i16, coeffs: *mut i32, tx_size: c_int, tx_type: c_int, bit_depth: c_int) -> c_int;
  pub fn r1c_fwd_txfm_batch(ctx: *mut r1c_ctx, tx_size: c_int, tx_type: c_int, bit_depth: c_int, residual: *const i16, coeffs: *mut i32, n: i64) -> c_int;
  pub fn r1c_fwd_txfm_batch_dev(ctx: *mut r1c_ctx, tx_size: c_int, tx_type: c_int, bit_depth: c_int, d_residual: *const i16, d_coeffs: *mut i32, n: i64) -> c_int;
  pub fn r1c_diff_fwd_txfm_batch(ctx: *mut r1c_ctx, input: *const r1c_plane, pred: *const r1c_plane, jobs: *const r1c_tx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, coeffs: *mut i32) -> c_int;
  pub fn r1c_diff_fwd_txfm_batch_dev(ctx: *mut r1c_ctx, input: *const r1c_plane, pred: *const r1c_plane, d_jobs: *const r1c_tx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, d_coeffs: *mut i32) -> c_int;
  pub fn r1c_inv_txfm_valid(tx_size: c_int, tx_type: c_int) -> c_int;
  pub fn r1c_inv_txfm_add_batch(ctx: *mut r1c_ctx, dst: *mut r1c_plane, jobs: *const r1c_itx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, rcoeffs: *const i32) -> c_int;
  pub fn r1c_inv_txfm_add_batch_dev(ctx: *mut r1c_ctx, dst: *mut r1c_plane, d_jobs: *const r1c_itx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, d_rcoeffs: *const i32) -> c_int;
  pub fn r1c_dequant_inv_txfm_add_batch(ctx: *mut r1c_ctx, dst: *mut r1c_plane, jobs: *const r1c_itx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, qcoeffs: *const i32, dc_quant: u32, ac_quant: u32) -> c_int;
  pub fn r1c_dequant_inv_txfm_add_batch_dev(ctx: *mut r1c_ctx, dst: *mut r1c_plane, d_jobs: *const r1c_itx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, d_qcoeffs: *const i32, dc_quant: u32, ac_quant: u32) -> c_int;
  pub fn r1c_log_tx_scale(tx_size: c_int) -> c_int;
  pub fn r1c_quant_ctx_update(qc: *mut r1c_quant_ctx, tx_size: c_int, is_intra: c_int, dc_quant: u32, ac_quant: u32) -> c_int;
  pub fn r1c_quantize_batch(ctx: *mut r1c_ctx, qc: *const r1c_quant_ctx, tx_size: c_int, scan: *const u16, coeffs: *const i32, qcoeffs: *mut i32, n: i64) -> c_int;
  pub fn r1c_dequantize_batch(ctx: *mut r1c_ctx, tx_size: c_int, dc_quant: u32, ac_quant: u32, qcoeffs: *const i32, rcoeffs: *mut i32, n: i64) -> c_int;
  pub fn r1c_quantize_batch_dev(ctx: *mut r1c_ctx, qc: *const r1c_quant_ctx, tx_size: c_int, d_scan: *const u16, d_coeffs: *const i32, d_qcoeffs: *mut i32, d_rcoeffs: *mut i32, d_tx_dist: *mut u64, n: i64) -> c_int;
  pub fn r1c_dequantize_batch_dev(ctx: *mut r1c_ctx, tx_size: c_int, dc_quant: u32, ac_quant: u32, d_qcoeffs: *const i32, d_rcoeffs: *mut i32, n: i64) -> c_int;
  pub fn r1c_encode_tx_block_valid(tx_size: c_int, tx_type: c_int) -> c_int;
  pub fn r1c_encode_tx_block_batch(ctx: *mut r1c_ctx, input: *const r1c_plane, rec: *mut r1c_plane, jobs: *const r1c_tx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, qc: *const r1c_quant_ctx, scan: *const u16, reconstruct: c_int, qcoeffs: *mut i32, tx_dist: *mut u64) -> c_int;
  pub fn r1c_encode_tx_block_batch_dev(ctx: *mut r1c_ctx, input: *const r1c_plane, rec: *mut r1c_plane, d_jobs: *const r1c_tx_job, n: i64, tx_size: c_int, tx_type: c_int, bit_depth: c_int, qc: *const r1c_quant_ctx, d_scan: *const u16, reconstruct: c_int, d_qcoeffs: *mut i32, d_tx_dist: *mut u64) -> c_int;
  pub fn r1c_mv_range(w_in_b: i64, h_in_b: i64, bo_x: i64, bo_y: i64, blk_w: c_int, blk_h: c_int, out: *mut i64);
  pub fn r1c_mv_rate(a: r1c_mv, b: r1c_mv, allow_high_precision_mv: c_int) -> u32;
  pub fn r1c_full_search_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_fs_job, n: i64, bit_depth: c_int, results: *mut r1c_me_result) -> c_int;
  pub fn r1c_full_search_batch_dev(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, d_jobs: *const r1c_fs_job, n: i64, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_sb_result_offset(bsize: c_int) -> c_int;
  pub fn r1c_sb_result_count(bsize: c_int) -> c_int;
  pub fn r1c_me_sb_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, params: *const r1c_sb_params, jobs: *const r1c_sb_job, n: i64, bit_depth: c_int, results: *mut r1c_me_result) -> c_int;
  pub fn r1c_me_sb_batch_dev(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, params: *const r1c_sb_params, d_jobs: *const r1c_sb_job, n: i64, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_me_sb_batch_dev_noalloc(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, params: *const r1c_sb_params, d_jobs: *const r1c_sb_job, n: i64, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_estimate_motion_ss4_batch(ctx: *mut r1c_ctx, org_qres: *const r1c_plane, ref_qres: *const r1c_plane, w_in_b: c_int, h_in_b: c_int, jobs: *const r1c_block_job, n: i64, me_range_scale: c_int, me_lambda: f64, allow_high_precision_mv: c_int, bit_depth: c_int, out: *mut r1c_mv) -> c_int;
  pub fn r1c_estimate_motion_ss2_batch(ctx: *mut r1c_ctx, org_hres: *const r1c_plane, ref_hres: *const r1c_plane, w_in_b: c_int, h_in_b: c_int, jobs: *const r1c_block_job, n: i64, pmvs: *const r1c_mv, pmv_valid: *const u8, me_lambda: f64, allow_high_precision_mv: c_int, bit_depth: c_int, out: *mut r1c_mv) -> c_int;
  pub fn r1c_estimate_motion_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, w_in_b: c_int, h_in_b: c_int, jobs: *const r1c_block_job, n: i64, cmvs: *const r1c_mv, cmv_valid: *const u8, me_lambda: f64, allow_high_precision_mv: c_int, bit_depth: c_int, out: *mut r1c_mv) -> c_int;
  pub fn r1c_save_block_motion_batch(ctx: *mut r1c_ctx, field: *mut r1c_mv, cols: c_int, rows: c_int, jobs: *const r1c_block_job, mvs: *const r1c_mv, n: i64) -> c_int;
  pub fn r1c_save_block_motion_batch_dev(ctx: *mut r1c_ctx, d_field: *mut r1c_mv, cols: c_int, rows: c_int, d_jobs: *const r1c_block_job, d_mvs: *const r1c_mv, n: i64) -> c_int;
  pub fn r1c_subset_predictors_batch(ctx: *mut r1c_ctx, field: *const r1c_mv, prev_field: *const r1c_mv, cols: c_int, rows: c_int, jobs: *const r1c_block_job, n: i64, cmvs: *const r1c_mv, n_cmvs: *const u8, predictors: *mut r1c_mv, counts: *mut i32) -> c_int;
  pub fn r1c_subset_predictors_batch_dev(ctx: *mut r1c_ctx, d_field: *const r1c_mv, d_prev_field: *const r1c_mv, cols: c_int, rows: c_int, d_jobs: *const r1c_block_job, n: i64, d_cmvs: *const r1c_mv, d_n_cmvs: *const u8, d_predictors: *mut r1c_mv, d_counts: *mut i32) -> c_int;
  pub fn r1c_subpel_cost_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_subpel_job, n: i64, cand: *const r1c_mv, n_cand: c_int, bit_depth: c_int, costs: *mut u64) -> c_int;
  pub fn r1c_subpel_refine_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_subpel_job, n: i64, bit_depth: c_int, results: *mut r1c_me_result) -> c_int;
  pub fn r1c_subpel_refine_batch_dev(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, d_jobs: *const r1c_subpel_job, n: i64, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_diamond_search_batch(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, jobs: *const r1c_subpel_job, n: i64, subpixel: c_int, predictors: *const r1c_mv, n_pred: c_int, bit_depth: c_int, results: *mut r1c_me_result) -> c_int;
  pub fn r1c_diamond_search_batch_dev(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, d_jobs: *const r1c_subpel_job, n: i64, subpixel: c_int, allow_high_precision_mv: c_int, d_predictors: *const r1c_mv, n_pred: c_int, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_frame_me_count(p: *const r1c_frame_me_params) -> i64;
  pub fn r1c_frame_me_offset(p: *const r1c_frame_me_params, bsize: c_int, cols: *mut i32, rows: *mut i32) -> i64;
  pub fn r1c_frame_me(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, p: *const r1c_frame_me_params, bit_depth: c_int, results: *mut r1c_me_result) -> c_int;
  pub fn r1c_frame_me_dev(ctx: *mut r1c_ctx, org: *const r1c_plane, r#ref: *const r1c_plane, p: *const r1c_frame_me_params, bit_depth: c_int, d_results: *mut r1c_me_result) -> c_int;
  pub fn r1c_frame_me_shard_count(ctx: *mut r1c_ctx, p: *const r1c_frame_me_params) -> i64;
  pub fn r1c_frame_me_compact_dev(ctx: *mut r1c_ctx, p: *const r1c_frame_me_params, d_results: *const r1c_me_result, d_compact: *mut r1c_me_result) -> c_int;
  pub fn r1c_frame_me_window(ctx: *mut r1c_ctx, pairs: *const r1c_frame_pair, n: i64, p: *const r1c_frame_me_params, bit_depth: c_int) -> c_int;
  pub fn r1c_frame_me_window_dev(ctx: *mut r1c_ctx, pairs: *const r1c_frame_pair, n: i64, p: *const r1c_frame_me_params, bit_depth: c_int) -> c_int;
  pub fn r1c_plane_upload_async(ctx: *mut r1c_ctx, plane: *mut r1c_plane, host_data: *const c_void) -> c_int;
  pub fn r1c_plane_upload_rows_async(ctx: *mut r1c_ctx, plane: *mut r1c_plane, host_data: *const c_void, row_lo: i64, row_hi: i64) -> c_int;
  pub fn r1c_plane_upload_rect_async(ctx: *mut r1c_ctx, plane: *mut r1c_plane, host_data: *const c_void, row_lo: i64, row_hi: i64, col_lo: i64, col_hi: i64) -> c_int;
  pub fn r1c_importance_propagate(ctx: *mut r1c_ctx, frame: *const r1c_plane, r#ref: *const r1c_plane, mvs: *const r1c_mv, mv_cols: c_int, mv_rows: c_int, intra_costs: *const u32, importances: *const f32, n_refs: c_int, w_in_imp_b: c_int, h_in_imp_b: c_int, bit_depth: c_int, ref_importances: *mut f32) -> c_int;
  pub fn r1c_importance_propagate_dev(ctx: *mut r1c_ctx, frame: *const r1c_plane, r#ref: *const r1c_plane, d_mvs: *const r1c_mv, mv_cols: c_int, mv_rows: c_int, d_intra_costs: *const u32, d_importances: *const f32, n_refs: c_int, w_in_imp_b: c_int, h_in_imp_b: c_int, bit_depth: c_int, d_ref_importances: *mut f32) -> c_int;
  pub fn r1c_importance_finalize(ctx: *mut r1c_ctx, intra_costs: *const u32, importances: *mut f32, n: i64) -> c_int;
  pub fn r1c_importance_finalize_dev(ctx: *mut r1c_ctx, d_intra_costs: *const u32, d_importances: *mut f32, n: i64) -> c_int;
}
