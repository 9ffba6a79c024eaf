/*
 * r1o.h -- CPU ORACLE for the rav1e hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of the reference's scalar ground-truth
 * functions (the `_ref` functions that rav1e's own `check_asm` feature
 * asserts every kernel against).  It is NOT part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load it.  The product (rav1e_b200/csrc) never links it.
 *
 * Parity pinning status (see DESIGN.md "Oracle"):
 *   SAD / SATD   : PINNED by the 44 golden values of src/dist.rs:379-402,437-460
 *   plane pad    : PINNED by the golden of src/frame/plane.rs:709-752
 *   fwd / inv TX : PINNED against the reference's own functions, whose Rust
 *                  text is transliterated and run at test time (tests/ref_fwd_tx.py,
 *                  tests/ref_inv_tx.py), and by golden vectors made that way
 *   MC, ME search, RDO distortion, quantize state machine, importances :
 *                  parity UNPINNED (the reference holds no golden vectors for
 *                  them and cannot be compiled here: no rustc / nasm).  They
 *                  are literal restatements cross-checked by independent
 *                  properties (tests/test_oracle_*.py).
 *
 * All strides are in PIXELS (the Rust API convention, plane_cfg.stride).
 * px_bytes is 1 (u8, bit_depth 8) or 2 (u16, bit_depth 10/12).
 */
#ifndef R1O_H
#define R1O_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums: numeric values equal the Rust declaration order ------------ */
/* BlockSize, src/partition.rs:116-140 */
enum {
  R1O_BLOCK_4X4, R1O_BLOCK_4X8, R1O_BLOCK_8X4, R1O_BLOCK_8X8, R1O_BLOCK_8X16,
  R1O_BLOCK_16X8, R1O_BLOCK_16X16, R1O_BLOCK_16X32, R1O_BLOCK_32X16,
  R1O_BLOCK_32X32, R1O_BLOCK_32X64, R1O_BLOCK_64X32, R1O_BLOCK_64X64,
  R1O_BLOCK_64X128, R1O_BLOCK_128X64, R1O_BLOCK_128X128, R1O_BLOCK_4X16,
  R1O_BLOCK_16X4, R1O_BLOCK_8X32, R1O_BLOCK_32X8, R1O_BLOCK_16X64,
  R1O_BLOCK_64X16, R1O_BLOCK_SIZES_ALL
};
/* FilterMode, src/mc.rs:61-66 */
enum { R1O_REGULAR = 0, R1O_SMOOTH = 1, R1O_SHARP = 2, R1O_BILINEAR = 3 };
/* TxSize, src/transform/mod.rs:225-247 ; TxType :123-140 */
enum {
  R1O_TX_4X4, R1O_TX_8X8, R1O_TX_16X16, R1O_TX_32X32, R1O_TX_64X64,
  R1O_TX_4X8, R1O_TX_8X4, R1O_TX_8X16, R1O_TX_16X8, R1O_TX_16X32,
  R1O_TX_32X16, R1O_TX_32X64, R1O_TX_64X32, R1O_TX_4X16, R1O_TX_16X4,
  R1O_TX_8X32, R1O_TX_32X8, R1O_TX_16X64, R1O_TX_64X16, R1O_TX_SIZES_ALL
};
enum {
  R1O_DCT_DCT, R1O_ADST_DCT, R1O_DCT_ADST, R1O_ADST_ADST, R1O_FLIPADST_DCT,
  R1O_DCT_FLIPADST, R1O_FLIPADST_FLIPADST, R1O_ADST_FLIPADST,
  R1O_FLIPADST_ADST, R1O_IDTX, R1O_V_DCT, R1O_H_DCT, R1O_V_ADST, R1O_H_ADST,
  R1O_V_FLIPADST, R1O_H_FLIPADST, R1O_TX_TYPES
};

int r1o_block_width(int bsize);
int r1o_block_height(int bsize);
int r1o_tx_width(int tx_size);
int r1o_tx_height(int tx_size);

/* ---- plane geometry: src/frame/plane.rs:22-47,215-244 ------------------- */
typedef struct {
  int64_t stride, alloc_height, width, height;
  int64_t xdec, ydec, xpad, ypad, xorigin, yorigin;
} r1o_plane_cfg;

void r1o_plane_cfg_new(r1o_plane_cfg *cfg, int64_t width, int64_t height,
                       int64_t xdec, int64_t ydec, int64_t xpad, int64_t ypad,
                       int px_bytes);
/* Plane::pad(w,h), src/frame/plane.rs:269-314.  data = whole allocation. */
void r1o_plane_pad(void *data, const r1o_plane_cfg *cfg, int64_t w, int64_t h,
                   int px_bytes);
/* Plane::downsample_from, src/frame/plane.rs:399-423.  0 ok, -1 size error */
int r1o_plane_downsample(void *dst, const r1o_plane_cfg *dcfg, const void *src,
                         const r1o_plane_cfg *scfg, int px_bytes);

/* ---- distortion: src/dist.rs:25-46 (SAD), :197-328 (SATD) --------------- */
uint32_t r1o_sad(const void *org, ptrdiff_t org_stride, const void *ref,
                 ptrdiff_t ref_stride, int w, int h, int px_bytes);
uint32_t r1o_satd(const void *org, ptrdiff_t org_stride, const void *ref,
                  ptrdiff_t ref_stride, int w, int h, int px_bytes);
/* the fork's generated kernels: i16 lanes for every pixel type (build/kernel/gen/dist.rs:171-350, :553-556) */
uint32_t r1o_satd_i16(const void *org, ptrdiff_t org_stride, const void *ref,
                      ptrdiff_t ref_stride, int w, int h, int px_bytes);

/* ---- motion compensation: src/mc.rs:201-408 ----------------------------- */
/* src points at the block origin (already integer-offset); filters read
 * 3 px left/up and 4 px right/down of it. */
void r1o_put_8tap(void *dst, ptrdiff_t dst_stride, const void *src,
                  ptrdiff_t src_stride, int w, int h, int col_frac,
                  int row_frac, int mode_x, int mode_y, int bit_depth,
                  int px_bytes);
void r1o_prep_8tap(int16_t *tmp, const void *src, ptrdiff_t src_stride, int w,
                   int h, int col_frac, int row_frac, int mode_x, int mode_y,
                   int bit_depth, int px_bytes);
void r1o_mc_avg(void *dst, ptrdiff_t dst_stride, const int16_t *tmp1,
                const int16_t *tmp2, int w, int h, int bit_depth,
                int px_bytes);
/* get_params of predict_inter, src/predict.rs:267-284: MV (1/8 luma pel) ->
 * clamped source origin (plane coords, may be negative) + 1/16 fracs. */
void r1o_mv_params(const r1o_plane_cfg *rec_cfg, int64_t po_x, int64_t po_y,
                   int mv_row, int mv_col, int *row_frac, int *col_frac,
                   int64_t *src_x, int64_t *src_y);

/* ---- inverse transform: src/transform/inverse.rs (r1o_itx.c) ------------- */
/* type1d: 0 DCT, 1 ADST, 2 FLIPADST, 3 identity; -1 where the reference has no function */
int r1o_inv_txfm1d(int type1d, int n, const int32_t *in, int32_t *out, int range);
int r1o_inv_txfm_valid(int tx_size, int tx_type);
/* coeffs: min(W,32)*min(H,32), row-major, stride min(W,32); residual: W*H COLUMN-major */
int r1o_inv_txfm2d(const int32_t *coeffs, int32_t *residual, int tx_size, int tx_type, int bd);
/* dst[r*stride + c] = clamp(dst + residual, 0, 2^bd - 1); stride in pixels */
int r1o_inv_txfm_add(const int32_t *coeffs, void *dst, ptrdiff_t stride, int px_bytes, int tx_size,
                     int tx_type, int bd);
/* n non-overlapping blocks: block i reads coeffs + i*coded, lands at (xy[2i], xy[2i+1]) of dst0 */
int r1o_inv_txfm_add_batch(const int32_t *coeffs, void *dst0, ptrdiff_t stride, int px_bytes,
                           const int32_t *xy, int64_t n, int tx_size, int tx_type, int bd);

/* ---- forward transform: src/transform/forward.rs ------------------------ */
/* 1-D kernels on n i32 values (n = 4..64); type: 0 DCT, 1 ADST, 2 IDTX.
 * returns -1 when the reference has `unimplemented!` for (type,n). */
int r1o_fwd_txfm1d(int type1d, int n, const int32_t *in, int32_t *out);
/* FwdTxfm2D::fht, forward.rs:1804-1899.  residual dense W*H (stride W) ->
 * coeffs dense W*H.  returns -1 for (size,type) pairs the reference cannot
 * instantiate.  Row-FLIPPED types use the intended semantic (column c is
 * written to W-1-c); the fork's SIMD branch for it is UB and unreachable
 * from the encoder. */
int r1o_fht(const int16_t *residual, int32_t *coeffs, int tx_size, int tx_type,
            int bit_depth);
int r1o_fht_valid(int tx_size, int tx_type);
/* diff, src/encoder.rs:1045-1059 */
void r1o_diff(int16_t *dst, const void *src1, ptrdiff_t stride1,
              const void *src2, ptrdiff_t stride2, int w, int h, int px_bytes);

/* ---- motion search: src/me.rs ------------------------------------------- */
typedef struct { int16_t row, col; } r1o_mv;
uint32_t r1o_mv_rate(r1o_mv a, r1o_mv b, int allow_hp); /* me.rs:1005-1021 */
/* get_mv_range, me.rs:64-80 : out = {mvx_min, mvx_max, mvy_min, mvy_max} */
void r1o_mv_range(int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y,
                  int blk_w, int blk_h, int64_t out[4]);
/* full_search, me.rs:943-990.  org/ref = pointers to plane coordinate (0,0)
 * (i.e. data + yorigin*stride + xorigin); po = block position; window
 * [x_lo..x_hi]x[y_lo..y_hi] step `step`.  best_mv / lowest_cost are in/out.
 * returns -1 if the search region leaves the allocated plane (the Rust
 * PlaneRegion::new asserts, src/tiling/plane_region.rs:139-142). */
int r1o_full_search(const void *org0, const r1o_plane_cfg *org_cfg,
                    const void *ref0, const r1o_plane_cfg *ref_cfg,
                    int64_t x_lo, int64_t x_hi, int64_t y_lo, int64_t y_hi,
                    int blk_w, int blk_h, int64_t po_x, int64_t po_y, int step,
                    uint32_t lambda, const r1o_mv pmv[2], int allow_hp,
                    int px_bytes, r1o_mv *best_mv, uint64_t *lowest_cost);
/* one sub-pel candidate: predict_inter (put_8tap REGULAR/REGULAR into a
 * blk-sized temp) + SAD|SATD + rate  (me.rs:787-856, 858-941).
 * returns UINT64_MAX if the MV is outside [mvx_min..mvx_max]x[...]. */
uint64_t r1o_subpel_cost(const void *org0, const r1o_plane_cfg *org_cfg,
                         const void *ref0, const r1o_plane_cfg *ref_cfg,
                         int64_t po_x, int64_t po_y, int blk_w, int blk_h,
                         r1o_mv cand, const r1o_mv pmv[2], uint32_t lambda,
                         int use_satd, int allow_hp, const int64_t range[4],
                         int bit_depth, int px_bytes);

/* ---- RDO distortion (r1o_rdo.c; src/rdo.rs:219-335,477-530, src/encoder.rs:1210-1224) ---- */
double r1o_distortion_bias(const float *importances, int64_t w_in_imp_b, int64_t w_in_b, int64_t h_in_b,
                           int64_t bo_x, int64_t bo_y, int64_t width_mi, int64_t height_mi);
uint64_t r1o_sse_wxh(const void *src1, int64_t stride1, int64_t x1, int64_t y1, const void *src2, int64_t stride2,
                     int64_t x2, int64_t y2, int64_t w, int64_t h, int xdec, int ydec, int px_bytes,
                     const float *importances, int64_t w_in_imp_b, int64_t w_in_b, int64_t h_in_b);
uint64_t r1o_cdef_dist_wxh(const void *src1, int64_t stride1, int64_t x1, int64_t y1, const void *src2,
                           int64_t stride2, int64_t x2, int64_t y2, int64_t w, int64_t h, int bit_depth,
                           int px_bytes, const float *importances, int64_t w_in_imp_b, int64_t w_in_b,
                           int64_t h_in_b);
uint64_t r1o_tx_dist(const int32_t *coeffs, const int32_t *rcoeffs, int tx_size);

/* ---- quantize / dequantize (r1o_quant.c), src/quantize.rs ---------------- */
/* QuantizationContext, src/quantize.rs:111-121 */
typedef struct {
  uint32_t log_tx_scale, dc_quant, ac_quant;
  int32_t dc_offset, ac_offset_eob, ac_offset0, ac_offset1;
  uint32_t dc_mul_add[3], ac_mul_add[3];
} r1o_quant_ctx;
int r1o_log_tx_scale(int tx_size);
void r1o_divu_gen(uint32_t d, uint32_t out[3]);
int32_t r1o_divu_pair(int32_t x, const uint32_t d[3]);
void r1o_quant_ctx_update(r1o_quant_ctx *qc, int tx_size, int is_intra, uint32_t dc_quant, uint32_t ac_quant);
void r1o_quantize(const r1o_quant_ctx *qc, const uint16_t *scan, int scan_len, const int32_t *coeffs, int32_t *qcoeffs);
void r1o_dequantize(uint32_t dc_quant, uint32_t ac_quant, int log_tx_scale, const int32_t *qcoeffs, int32_t *rcoeffs,
                    int len);
void r1o_quantize_batch(const r1o_quant_ctx *qc, int tx_size, const uint16_t *scan, const int32_t *coeffs,
                        int32_t *qcoeffs, int32_t *rcoeffs, int64_t n);

/* ---- block importances (r1o_imp.c), src/api/internal.rs:823-1075 ---------- */
void r1o_importance_propagate(const void *frame0, const r1o_plane_cfg *fcfg, const void *ref0, const r1o_plane_cfg *rcfg,
                              const int16_t *mvs, int mv_cols, const uint32_t *intra_costs, const float *importances,
                              int n_refs, int w, int h, int bit_depth, int px_bytes, float *ref_importances);
void r1o_importance_finalize(const uint32_t *intra_costs, float *importances, int64_t n);

#ifdef __cplusplus
}
#endif
/* benches/dist.rs recipe: ns per get_sad / get_satd call, one thread (bench.py configs[0] line) */
double r1o_bench_dist(int use_satd, const void *org, ptrdiff_t os, const void *ref, ptrdiff_t rs, int w, int h, int pb,
                      int64_t iters);
#endif /* R1O_H */
