/*
 * rav1e_cuda.h -- C ABI of librav1e_cuda.so: the B200 (sm_100a) implementation
 * of rav1e's data-parallel inner loop (distortion, sub-pel motion
 * compensation, forward transforms, motion search).
 *
 * Boundary it replaces.  The reference (geobacter-rs/rav1e) has no FFI for
 * this path; its closest equivalent -- and the pattern followed here -- is
 * the C-ABI function-pointer boundary of its asm back-ends:
 *     SadFn / SatdFn      src/asm/x86/dist.rs:16-32
 *     PutFn / PutHBDFn    src/asm/x86/mc.rs:17-38
 *     PrepFn / PrepHBDFn  src/asm/x86/mc.rs:40-59
 *     AvgFn / AvgHBDFn    src/asm/x86/mc.rs:61-78
 * plus the pure-Rust entry points that have no asm today:
 *     forward_transform   src/transform/mod.rs:556-566
 *     full_search         src/me.rs:943-990
 *     telescopic_subpel_search  src/me.rs:858-941
 * Conventions kept from that boundary: the caller owns every buffer, the
 * callee never allocates caller-visible memory, preconditions are checked
 * (the Rust side `assert!`s them, here they return R1C_ERR_*), functions are
 * re-entrant per context (one r1c_ctx per rayon worker / tile).
 *
 * Conventions of this header
 *   - every entry point returns 0 (R1C_OK) or a negative R1C_ERR_* code and
 *     never aborts; r1c_last_error() gives the text for the last failure.
 *   - enum values equal the Rust declaration order (BlockSize
 *     src/partition.rs:116-140, TxSize / TxType src/transform/mod.rs:225-247,
 *     123-140, FilterMode src/mc.rs:61-66).
 *   - coordinates are plane coordinates (0,0 = first visible pixel; negative
 *     values address the padding), strides are in PIXELS as in PlaneConfig.
 *   - `*_batch` entry points take HOST job / result arrays and do the
 *     host<->device copies themselves (the end-to-end path);
 *     `*_batch_dev` take DEVICE job / result arrays, only enqueue work on the
 *     context's stream and return (r1c_sync() to wait).
 *   - there is no CPU fallback: without a CUDA device r1c_create() fails.
 */
#ifndef RAV1E_CUDA_H
#define RAV1E_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define R1C_OK 0
#define R1C_ERR_ARG (-1)         /* bad argument / unsupported combination   */
#define R1C_ERR_CUDA (-2)        /* CUDA runtime error                       */
#define R1C_ERR_BOUNDS (-3)      /* region leaves the allocated plane        */
#define R1C_ERR_UNIMPLEMENTED (-4) /* `unimplemented!` in the reference too  */
#define R1C_ERR_NOMEM (-5)

/* BlockSize, src/partition.rs:116-140 */
typedef enum {
  R1C_BLOCK_4X4, R1C_BLOCK_4X8, R1C_BLOCK_8X4, R1C_BLOCK_8X8, R1C_BLOCK_8X16,
  R1C_BLOCK_16X8, R1C_BLOCK_16X16, R1C_BLOCK_16X32, R1C_BLOCK_32X16,
  R1C_BLOCK_32X32, R1C_BLOCK_32X64, R1C_BLOCK_64X32, R1C_BLOCK_64X64,
  R1C_BLOCK_64X128, R1C_BLOCK_128X64, R1C_BLOCK_128X128, R1C_BLOCK_4X16,
  R1C_BLOCK_16X4, R1C_BLOCK_8X32, R1C_BLOCK_32X8, R1C_BLOCK_16X64,
  R1C_BLOCK_64X16, R1C_BLOCK_SIZES_ALL
} r1c_block_size;

/* FilterMode, src/mc.rs:61-66 */
typedef enum { R1C_REGULAR = 0, R1C_SMOOTH = 1, R1C_SHARP = 2, R1C_BILINEAR = 3 } r1c_filter_mode;

/* TxSize, src/transform/mod.rs:225-247 */
typedef enum {
  R1C_TX_4X4, R1C_TX_8X8, R1C_TX_16X16, R1C_TX_32X32, R1C_TX_64X64,
  R1C_TX_4X8, R1C_TX_8X4, R1C_TX_8X16, R1C_TX_16X8, R1C_TX_16X32,
  R1C_TX_32X16, R1C_TX_32X64, R1C_TX_64X32, R1C_TX_4X16, R1C_TX_16X4,
  R1C_TX_8X32, R1C_TX_32X8, R1C_TX_16X64, R1C_TX_64X16, R1C_TX_SIZES_ALL
} r1c_tx_size;

/* TxType, src/transform/mod.rs:123-140 */
typedef enum {
  R1C_DCT_DCT, R1C_ADST_DCT, R1C_DCT_ADST, R1C_ADST_ADST, R1C_FLIPADST_DCT,
  R1C_DCT_FLIPADST, R1C_FLIPADST_FLIPADST, R1C_ADST_FLIPADST,
  R1C_FLIPADST_ADST, R1C_IDTX, R1C_V_DCT, R1C_H_DCT, R1C_V_ADST, R1C_H_ADST,
  R1C_V_FLIPADST, R1C_H_FLIPADST, R1C_TX_TYPES
} r1c_tx_type;

int r1c_block_width(int bsize);   /* -1 if out of range */
int r1c_block_height(int bsize);
int r1c_tx_width(int tx_size);
int r1c_tx_height(int tx_size);
/* 1 if the reference can instantiate (tx_size, tx_type): DCT 4..64,
 * ADST/FLIPADST 4..16, identity 4..32 (src/transform/forward.rs:1745-1768) */
int r1c_fwd_txfm_valid(int tx_size, int tx_type);

/* ------------------------------------------------------------------------
 * Context: one device + one stream + scratch.  Not shared between threads.
 * (replaces nothing in the reference: the asm boundary is stateless; state is
 * needed here for the device, the stream and pinned staging buffers.)
 * ---------------------------------------------------------------------- */
typedef struct r1c_ctx r1c_ctx;
r1c_ctx *r1c_create(int device);          /* NULL when there is no usable CUDA device */
void r1c_destroy(r1c_ctx *ctx);
int r1c_sync(r1c_ctx *ctx);               /* wait for everything enqueued on the ctx stream */
const char *r1c_last_error(const r1c_ctx *ctx);
void *r1c_stream(r1c_ctx *ctx);           /* the cudaStream_t, for event timing */
int r1c_device(const r1c_ctx *ctx);
/* number of kernels this context has launched so far (bench.py gpu_launches) */
int64_t r1c_launch_count(const r1c_ctx *ctx);
/* Device-side timing of one kernel class, for the roofline report: while
 * enabled, the launches of that class are bracketed by CUDA events on the
 * context stream.  r1c_profile_read waits for the stream, returns the summed
 * duration and the number of launches since the last read, and resets. */
#define R1C_PROF_ME_SB 0      /* me_sb_kernel (superblock full search)        */
#define R1C_PROF_SUBPEL_MC 1  /* put_8tap of the sub-pel candidates           */
#define R1C_PROF_SUBPEL_COST 2 /* SAD / SATD of the predicted candidates      */
#define R1C_PROF_MC 3         /* put / prep / avg batch kernels               */
#define R1C_PROF_TX 4         /* forward transform kernels                    */
#define R1C_PROF_DIST 5       /* SAD / SATD / RDO distortion batch kernels    */
#define R1C_PROF_QUANT 6      /* quantize / dequantize kernels                */
#define R1C_PROF_ITX 7        /* inverse transform + add kernels              */
#define R1C_PROF_SLOTS 8
/* Diagnostics: measured rate of one integer instruction class on this context's device, in giga lane-operations per
 * second (dependency-free streams on every SM; a few milliseconds).  The INT roofline denominator bench.py reports
 * against (a measured peak, not the nominal one).  No counterpart in the reference. */
#define R1C_PEAK_IADD 0       /* IADD3: the scalar 32-bit lane rate */
#define R1C_PEAK_VABSDIFF4 1  /* VABSDIFF4 with accumulate: 8 byte-ops per lane-op */
#define R1C_PEAK_DP4A 2
#define R1C_PEAK_DP2A 3
#define R1C_PEAK_IMAD 4
#define R1C_PEAK_PRMT 5
int r1c_int_peak(r1c_ctx *ctx, int which, double *giga_lane_ops_per_s);
int r1c_profile(r1c_ctx *ctx, int enable);
int r1c_profile_read(r1c_ctx *ctx, int slot, double *ms, int64_t *launches);

/* ------------------------------------------------------------------------
 * Planes.  Mirrors Plane<T> / PlaneConfig (src/frame/plane.rs:22-47,154,
 * 215-244): a padded allocation of stride*alloc_height pixels, visible area
 * starting at (xorigin, yorigin).
 * ---------------------------------------------------------------------- */
typedef struct {
  int64_t stride, alloc_height, width, height;
  int64_t xdec, ydec, xpad, ypad, xorigin, yorigin;
} r1c_plane_cfg;

typedef struct r1c_plane r1c_plane;

/* Plane::new geometry (src/frame/plane.rs:215-244); px_bytes 1 (u8) or 2 (u16) */
int r1c_plane_cfg_new(r1c_plane_cfg *cfg, int64_t width, int64_t height, int64_t xdec,
                      int64_t ydec, int64_t xpad, int64_t ypad, int px_bytes);
/* device allocation, every pixel 128 like PlaneData::new (plane.rs:130-138) */
int r1c_plane_create(r1c_ctx *ctx, const r1c_plane_cfg *cfg, int px_bytes, r1c_plane **out);
/* adopt caller-owned DEVICE memory holding a whole allocation (not freed) */
int r1c_plane_wrap(r1c_ctx *ctx, void *device_data, const r1c_plane_cfg *cfg, int px_bytes,
                   r1c_plane **out);
void r1c_plane_destroy(r1c_plane *plane);
/* whole allocation (stride*alloc_height pixels) host -> device / device -> host */
int r1c_plane_upload(r1c_ctx *ctx, r1c_plane *plane, const void *host_data);
int r1c_plane_download(r1c_ctx *ctx, const r1c_plane *plane, void *host_data);
void *r1c_plane_device_ptr(const r1c_plane *plane);
const r1c_plane_cfg *r1c_plane_config(const r1c_plane *plane);
/* Plane::pad(w, h), src/frame/plane.rs:269-314, on the device */
int r1c_plane_pad(r1c_ctx *ctx, r1c_plane *plane, int64_t w, int64_t h);
/* Plane::downsample_from, src/frame/plane.rs:399-423, on the device */
int r1c_plane_downsample(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src);

/* ------------------------------------------------------------------------
 * Distortion: get_sad / get_satd, src/dist.rs:48-111,121-188 (semantics of
 * get_sad_ref :25-46 and get_satd_ref :197-328).
 * ---------------------------------------------------------------------- */
typedef struct {
  int32_t org_x, org_y; /* block origin in the `org` plane */
  int32_t ref_x, ref_y; /* block origin in the `ref` plane */
  int32_t bsize;        /* r1c_block_size */
} r1c_dist_job;

/* one block, blocking, result to host: signature-compatible with SadFn /
 * SatdFn (region pointers become plane + offset). */
int r1c_sad(r1c_ctx *ctx, const r1c_plane *org, int org_x, int org_y, const r1c_plane *ref,
            int ref_x, int ref_y, int bsize, int bit_depth, uint32_t *out);
int r1c_satd(r1c_ctx *ctx, const r1c_plane *org, int org_x, int org_y, const r1c_plane *ref,
             int ref_x, int ref_y, int bsize, int bit_depth, uint32_t *out);
/* n jobs -> n u32 */
int r1c_sad_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                  const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out);
int r1c_satd_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                   const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out);
/* `use_satd` of the device entry: R1C_DIST_SAD, R1C_DIST_SATD, or R1C_DIST_SATD_I16 */
#define R1C_DIST_SAD 0
#define R1C_DIST_SATD 1
#define R1C_DIST_SATD_I16 2
int r1c_dist_batch_dev(r1c_ctx *ctx, int use_satd, const r1c_plane *org, const r1c_plane *ref,
                       const r1c_dist_job *d_jobs, int64_t n, int bit_depth, uint32_t *d_out);
/* The `compat_i16_wrap` switch.  r1c_satd* follow get_satd_ref (i32 arithmetic,
 * src/dist.rs:197-328), which is what the reference's check_asm asserts against.  The kernels
 * this fork GENERATES and ships (build/kernel/gen/dist.rs:171-350) compute in i16 lanes for every
 * pixel type (`let calc_prim = PrimType::I16`, :189-194, :252-257): identical for 8-bit input,
 * but 10- and 12-bit blocks of high amplitude wrap (a term of 2^15 or more after the doubling).  This entry reproduces THAT arithmetic
 * (wrapping i16 butterflies, abs().max() * 2 on the merged last stage, sign-extending cast to
 * u32, wrapping sum) so that results can be compared with a reference binary built from the
 * fork.  Not a fast path. */
int r1c_satd_i16_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                       const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out);

/* compute_lookahead_intra_costs, src/api/internal.rs:680-765: out[y*w_in_imp_b + x] = get_satd of the
 * 8x8 importance block (x, y) of the input luma plane against its DC_PRED prediction.  The
 * reference hands predict_intra the block's own rectangle as tile_rect, so the variant is always
 * NONE and the prediction is pred_dc_128 = 128 << (bit_depth - 8) for every block
 * (src/predict.rs:176-184,206-212,623-633); reproduced as is. */
int r1c_lookahead_intra_costs(r1c_ctx *ctx, const r1c_plane *input, int w_in_imp_b, int h_in_imp_b,
                              int bit_depth, uint32_t *out);
int r1c_lookahead_intra_costs_dev(r1c_ctx *ctx, const r1c_plane *input, int w_in_imp_b, int h_in_imp_b,
                                  int bit_depth, uint32_t *d_out);

/* ------------------------------------------------------------------------
 * Motion compensation: put_8tap / prep_8tap / mc_avg, src/mc.rs:410-706
 * (semantics of the `_ref` functions :213-408).
 * `src_x/src_y` is the block origin in the source plane AFTER the integer MV
 * offset and the clamp of predict_inter (src/predict.rs:267-284; see
 * r1c_mv_params); the filters read 3 px left/up and 4 px right/down of it.
 * ---------------------------------------------------------------------- */
typedef struct {
  int32_t dst_x, dst_y;       /* put: block origin in the dst plane; prep: unused */
  int32_t src_x, src_y;
  int16_t w, h;               /* 2..128; any size, not only BlockSize shapes */
  int8_t col_frac, row_frac;  /* 0..15, 1/16 px */
  int8_t mode_x, mode_y;      /* r1c_filter_mode */
  int32_t tmp_off;            /* prep: offset (in i16 elements) of the dense w*h output */
} r1c_mc_job;

typedef struct {
  int32_t dst_x, dst_y;
  int16_t w, h;
  int32_t tmp_off; /* offset (i16 elements) of the dense w*h block inside tmp1 and tmp2 */
} r1c_avg_job;

/* predict_inter's get_params (src/predict.rs:267-284): MV in 1/8 luma px ->
 * 1/16 fracs and the clamped source origin.  Pure host arithmetic. */
void r1c_mv_params(const r1c_plane_cfg *rec_cfg, int64_t po_x, int64_t po_y, int mv_row,
                   int mv_col, int *row_frac, int *col_frac, int64_t *src_x, int64_t *src_y);

int r1c_put_8tap_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src, const r1c_mc_job *jobs,
                       int64_t n, int bit_depth);
/* tmp: HOST i16 array of tmp_len elements, written at jobs[i].tmp_off */
int r1c_prep_8tap_batch(r1c_ctx *ctx, int16_t *tmp, int64_t tmp_len, const r1c_plane *src,
                        const r1c_mc_job *jobs, int64_t n, int bit_depth);
int r1c_mc_avg_batch(r1c_ctx *ctx, r1c_plane *dst, const int16_t *tmp1, const int16_t *tmp2,
                     int64_t tmp_len, const r1c_avg_job *jobs, int64_t n, int bit_depth);
/* device-resident variants (jobs, tmp on the device) */
int r1c_put_8tap_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src,
                           const r1c_mc_job *d_jobs, int64_t n, int bit_depth);
int r1c_prep_8tap_batch_dev(r1c_ctx *ctx, int16_t *d_tmp, const r1c_plane *src,
                            const r1c_mc_job *d_jobs, int64_t n, int bit_depth);
int r1c_mc_avg_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const int16_t *d_tmp1,
                         const int16_t *d_tmp2, const r1c_avg_job *d_jobs, int64_t n,
                         int bit_depth);

/* ------------------------------------------------------------------------
 * RDO distortion.
 *   r1c_sse_wxh_batch    sse_wxh, src/rdo.rs:286-335: SSE per importance block
 *                        (8x8 luma px, decimated with src1's plane) times the
 *                        f64 distortion bias, truncated to u64 per block, summed
 *   r1c_cdef_dist_batch  cdef_dist_wxh, src/rdo.rs:219-283 (w, h multiples of 8,
 *                        non-subsampled planes)
 *   bias                 compute_distortion_bias, src/rdo.rs:477-508, from
 *                        fi.block_importances (src/encoder.rs:719) at the frame
 *                        block offset of (src_x, src_y) in src1's plane;
 *                        block_importances == NULL gives bias 1.0 (raw sums)
 *   r1c_tx_dist_batch    tx-domain distortion, src/encoder.rs:1210-1224: block i
 *                        reads coeffs + i*W*H and rcoeffs + i*min(W,32)*min(H,32)
 * Results are Distortion values (u64), bit-identical to the reference's f64
 * arithmetic (no fused multiply-add, IEEE div / sqrt).
 * ---------------------------------------------------------------------- */
typedef struct {
  int32_t src_x, src_y; /* region origin in src1 (the input plane), >= 0 */
  int32_t rec_x, rec_y; /* region origin in src2 (the reconstruction)    */
  int32_t w, h;         /* pixels of these planes                        */
} r1c_rdo_job;
typedef struct {
  int32_t w_in_imp_b; /* fi.w_in_imp_b: row pitch of block_importances     */
  int32_t n_imp;      /* elements in block_importances                     */
  int32_t w_in_b;     /* fi.w_in_b, fi.h_in_b: frame size in 4x4 units     */
  int32_t h_in_b;
} r1c_bias_params;
int r1c_sse_wxh_batch(r1c_ctx *ctx, const r1c_plane *src1, const r1c_plane *src2,
                      const r1c_rdo_job *jobs, int64_t n, int bit_depth,
                      const float *block_importances, const r1c_bias_params *bp, uint64_t *out);
int r1c_cdef_dist_batch(r1c_ctx *ctx, const r1c_plane *src1, const r1c_plane *src2,
                        const r1c_rdo_job *jobs, int64_t n, int bit_depth,
                        const float *block_importances, const r1c_bias_params *bp, uint64_t *out);
/* kind 0 = sse_wxh, 1 = cdef_dist_wxh; every pointer is device memory except bp */
int r1c_rdo_dist_batch_dev(r1c_ctx *ctx, int kind, const r1c_plane *src1, const r1c_plane *src2,
                           const r1c_rdo_job *d_jobs, int64_t n, int bit_depth,
                           const float *d_block_importances, const r1c_bias_params *bp,
                           uint64_t *d_out);
int r1c_tx_dist_batch(r1c_ctx *ctx, const int32_t *coeffs, const int32_t *rcoeffs, int tx_size,
                      int64_t n, uint64_t *out);
int r1c_tx_dist_batch_dev(r1c_ctx *ctx, const int32_t *d_coeffs, const int32_t *d_rcoeffs,
                          int tx_size, int64_t n, uint64_t *d_out);

/* ------------------------------------------------------------------------
 * Forward transforms: forward_transform, src/transform/mod.rs:556-566 ->
 * FwdTxfm2D::fht, src/transform/forward.rs:1804-1899.
 * residual: dense W*H i16, row-major (stride W); coeffs: dense W*H i32,
 * coeffs[r*W + c].  Blocks of one call are stored back to back.
 * Row-FLIPPED types (DCT_FLIPADST, FLIPADST_FLIPADST, ADST_FLIPADST,
 * H_FLIPADST) use the intended semantic "column c is written to W-1-c"
 * (the fork's SIMD branch forward.rs:1858-1862 reads uninitialised memory
 * and is unreachable from the encoder, which never selects FLIPADST).
 * ---------------------------------------------------------------------- */
int r1c_forward_transform(r1c_ctx *ctx, const int16_t *residual, int32_t *coeffs, int tx_size,
                          int tx_type, int bit_depth); /* one block, host buffers */
int r1c_fwd_txfm_batch(r1c_ctx *ctx, int tx_size, int tx_type, int bit_depth,
                       const int16_t *residual, int32_t *coeffs, int64_t n);
int r1c_fwd_txfm_batch_dev(r1c_ctx *ctx, int tx_size, int tx_type, int bit_depth,
                           const int16_t *d_residual, int32_t *d_coeffs, int64_t n);
/* fused diff (src/encoder.rs:1045-1059) + fht: residual = input - pred taken
 * straight from two planes; block i -> coeffs + i*W*H */
typedef struct {
  int32_t in_x, in_y;     /* block origin in the input plane */
  int32_t pred_x, pred_y; /* block origin in the prediction plane */
} r1c_tx_job;
int r1c_diff_fwd_txfm_batch(r1c_ctx *ctx, const r1c_plane *input, const r1c_plane *pred,
                            const r1c_tx_job *jobs, int64_t n, int tx_size, int tx_type,
                            int bit_depth, int32_t *coeffs);
int r1c_diff_fwd_txfm_batch_dev(r1c_ctx *ctx, const r1c_plane *input, const r1c_plane *pred,
                                const r1c_tx_job *d_jobs, int64_t n, int tx_size, int tx_type,
                                int bit_depth, int32_t *d_coeffs);

/* ------------------------------------------------------------------------
 * Inverse transforms: inverse_transform_add, src/transform/mod.rs:568-580 ->
 * InvTxfm2D::inv_txfm2d_add, src/transform/inverse.rs:1939-2028 (inv_txfm2d :2061-2107, the 1-D
 * functions :35-1544).
 * rcoeffs: block i reads rcoeffs + i*coded, coded = min(W,32)*min(H,32), row-major with stride
 * min(W,32) -- what dequantize leaves (64-point transforms signal 32 coefficients per line).
 * Each block is transformed and ADDED to the W x H pixels at (jobs[i].x, jobs[i].y) of `dst`
 * (visible-area coordinates), clamped to [0, 2^bit_depth - 1].  Blocks of one call must not
 * overlap.  (size, type) pairs: every 1-D function INV_TXFM_FNS holds (:1610-1634; DCT 4..64, ADST /
 * FLIPADST 4..16, identity 4..32) -- the same 193 pairs as the forward side; FLIPADST is
 * av1_iflipadst* (the reversed ADST output, :57-60), which the fork's portable Rust path leaves
 * `unimplemented!` (:1596-1606) and only reaches through its generated kernels.
 * ---------------------------------------------------------------------- */
typedef struct {
  int32_t x, y; /* block origin in the destination plane */
} r1c_itx_job;
int r1c_inv_txfm_valid(int tx_size, int tx_type);
int r1c_inv_txfm_add_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *jobs, int64_t n,
                           int tx_size, int tx_type, int bit_depth, const int32_t *rcoeffs);
/* d_jobs and d_rcoeffs are device arrays; the caller guarantees the jobs stay inside the plane */
int r1c_inv_txfm_add_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *d_jobs, int64_t n,
                               int tx_size, int tx_type, int bit_depth, const int32_t *d_rcoeffs);
/* dequantize (src/quantize.rs:318-334) fused into the load of the row pass: the array holds
 * QUANTIZED coefficients, same layout; the pair of calls at src/encoder.rs:1192-1209 in one pass,
 * without the rcoeffs array in between */
int r1c_dequant_inv_txfm_add_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *jobs, int64_t n,
                                   int tx_size, int tx_type, int bit_depth, const int32_t *qcoeffs,
                                   uint32_t dc_quant, uint32_t ac_quant);
int r1c_dequant_inv_txfm_add_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *d_jobs,
                                       int64_t n, int tx_size, int tx_type, int bit_depth,
                                       const int32_t *d_qcoeffs, uint32_t dc_quant, uint32_t ac_quant);

/* ------------------------------------------------------------------------
 * Quantize / dequantize, src/quantize.rs.
 *   r1c_quant_ctx            QuantizationContext, src/quantize.rs:111-121
 *   r1c_quant_ctx_update     QuantizationContext::update, :212-255, from the VALUES
 *                            dc_q(qindex, dc_delta_q, bit_depth) / ac_q(..) (:41-61; the
 *                            qindex tables stay on the Rust side)
 *   r1c_quantize_batch       QuantizationContext::quantize, :257-316, T = i32
 *                            (the encoder's call, src/encoder.rs:1170)
 *   r1c_dequantize_batch     dequantize, :318-334
 * Layout = the encoder's (src/encoder.rs:1143-1155): block i reads coeffs + i*W*H and
 * writes qcoeffs / rcoeffs + i*coded, coded = min(W,32)*min(H,32) (av1_get_coded_tx_size);
 * `scan` = av1_scan_orders[tx_size][tx_type].scan, `coded` entries, scan[0] == 0, a
 * permutation of 0..coded-1 (host variants check it: R1C_ERR_ARG otherwise).
 * The fused device entry also writes the reconstruction coefficients and the tx-domain
 * distortion of src/encoder.rs:1210-1224 in the same pass (either may be NULL).
 * ---------------------------------------------------------------------- */
typedef struct {
  uint32_t log_tx_scale, dc_quant, ac_quant;
  int32_t dc_offset, ac_offset_eob, ac_offset0, ac_offset1;
  uint32_t dc_mul_add[3], ac_mul_add[3]; /* divu_gen, src/quantize.rs:123-138 */
} r1c_quant_ctx;
int r1c_log_tx_scale(int tx_size); /* get_log_tx_scale, src/quantize.rs:34-39 */
int r1c_quant_ctx_update(r1c_quant_ctx *qc, int tx_size, int is_intra, uint32_t dc_quant,
                         uint32_t ac_quant);
int r1c_quantize_batch(r1c_ctx *ctx, const r1c_quant_ctx *qc, int tx_size, const uint16_t *scan,
                       const int32_t *coeffs, int32_t *qcoeffs, int64_t n);
int r1c_dequantize_batch(r1c_ctx *ctx, int tx_size, uint32_t dc_quant, uint32_t ac_quant,
                         const int32_t *qcoeffs, int32_t *rcoeffs, int64_t n);
/* device arrays; d_rcoeffs and d_tx_dist are optional outputs */
int r1c_quantize_batch_dev(r1c_ctx *ctx, const r1c_quant_ctx *qc, int tx_size,
                           const uint16_t *d_scan, const int32_t *d_coeffs, int32_t *d_qcoeffs,
                           int32_t *d_rcoeffs, uint64_t *d_tx_dist, int64_t n);
int r1c_dequantize_batch_dev(r1c_ctx *ctx, int tx_size, uint32_t dc_quant, uint32_t ac_quant,
                             const int32_t *d_qcoeffs, int32_t *d_rcoeffs, int64_t n);

/* ------------------------------------------------------------------------
 * encode_tx_block, src/encoder.rs:1143-1224 -- the residual chain of one transform block as ONE
 * kernel: diff (:1045-1059, :1157-1163) -> forward_transform (:1165-1168) -> quantize (:1170) ->
 * dequantize (:1192-1200) -> inverse_transform_add (:1203-1208) -> tx-domain distortion
 * (:1210-1224).  The entropy coder (write_coeffs_lv_map, :1173-1190) stays on the Rust side: it
 * reads `qcoeffs`.  Intra prediction (:1096-1119) is not part of the call: `rec` must already
 * hold the block's prediction (what motion_compensate / predict_intra leave there), it is the
 * subtrahend of the diff and, when `reconstruct` is set, receives the reconstruction
 * (`!fi.use_tx_domain_distortion || need_recon_pixel`).  Blocks of one call are independent in
 * the reference only if no block's prediction reads another's reconstruction (inter blocks, or
 * intra blocks of one wave front): the jobs of one call must not overlap in `rec`.
 *   jobs[i]   (in_x, in_y) in `input`, (pred_x, pred_y) in `rec`
 *   qcoeffs   block i at qcoeffs + i*min(W,32)*min(H,32) (av1_get_coded_tx_size): all 19 TxSize.
 *             The five with a 64-point side follow this fork's two layouts (dense 64-wide
 *             forward rows, of which the quantizer and the inverse see the first coded-area
 *             words -- the inverse as 32-wide rows), like the separate entry points.
 *             r1c_encode_tx_block_valid() == 0 for the (size, type) pairs the reference leaves
 *             `unimplemented!`; the batch calls then return R1C_ERR_UNIMPLEMENTED
 *   tx_dist   optional (`rdo_type.needs_tx_dist()`): sum((coeff - rcoeff)^2) rounded and
 *             shifted by 2*(3 - log_tx_scale), before the bias / dist_scale factors of :1238-1240
 *   qc, scan  as for r1c_quantize_batch
 * Bit-exact with the chain r1c_diff_fwd_txfm_batch -> r1c_quantize_batch_dev ->
 * r1c_dequant_inv_txfm_add_batch.
 * ---------------------------------------------------------------------- */
int r1c_encode_tx_block_valid(int tx_size, int tx_type);
int r1c_encode_tx_block_batch(r1c_ctx *ctx, const r1c_plane *input, r1c_plane *rec,
                              const r1c_tx_job *jobs, int64_t n, int tx_size, int tx_type,
                              int bit_depth, const r1c_quant_ctx *qc, const uint16_t *scan,
                              int reconstruct, int32_t *qcoeffs, uint64_t *tx_dist);
/* d_jobs, d_scan, d_qcoeffs, d_tx_dist are device arrays; the caller guarantees the jobs stay
 * inside the planes and that d_scan is a permutation with d_scan[0] == 0 */
int r1c_encode_tx_block_batch_dev(r1c_ctx *ctx, const r1c_plane *input, r1c_plane *rec,
                                  const r1c_tx_job *d_jobs, int64_t n, int tx_size, int tx_type,
                                  int bit_depth, const r1c_quant_ctx *qc, const uint16_t *d_scan,
                                  int reconstruct, int32_t *d_qcoeffs, uint64_t *d_tx_dist);

/* ------------------------------------------------------------------------
 * Motion search, src/me.rs.
 * ---------------------------------------------------------------------- */
typedef struct { int16_t row, col; } r1c_mv; /* MotionVector, src/mc.rs:27-31; 1/8 px */

typedef struct {
  r1c_mv mv;
  uint32_t reserved;
  uint64_t cost; /* 256*dist + rate*lambda, src/me.rs:855,982; UINT64_MAX = none */
} r1c_me_result;

/* get_mv_range, src/me.rs:64-80: out = {mvx_min, mvx_max, mvy_min, mvy_max} */
void r1c_mv_range(int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w,
                  int blk_h, int64_t out[4]);
/* get_mv_rate, src/me.rs:1005-1021 */
uint32_t r1c_mv_rate(r1c_mv a, r1c_mv b, int allow_high_precision_mv);

/* full_search, src/me.rs:943-990: every (y, x) of [y_lo..y_hi] x [x_lo..x_hi]
 * with stride `step`, raster order, strict `<` (first minimum wins). */
typedef struct {
  int32_t po_x, po_y;              /* block origin (both planes) */
  int32_t bsize;                   /* r1c_block_size */
  int32_t x_lo, x_hi, y_lo, y_hi;  /* candidate positions in the ref plane */
  int32_t step;
  uint32_t lambda;
  r1c_mv pmv[2];
  int32_t allow_high_precision_mv;
  int32_t has_init; /* != 0: results[i] holds best_mv / lowest_cost on entry */
} r1c_fs_job;

int r1c_full_search_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                          const r1c_fs_job *jobs, int64_t n, int bit_depth,
                          r1c_me_result *results /* in/out */);
int r1c_full_search_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                              const r1c_fs_job *d_jobs, int64_t n, int bit_depth,
                              r1c_me_result *d_results);

/* Whole-superblock full search: ONE pass over the reference window yields the
 * result of full_search for EVERY block of every requested size inside a
 * 64x64 region (4x4 SADs are computed once per candidate and summed up to
 * the larger blocks).  Per block the window is the one
 * FullSearch::full_pixel_me builds (src/me.rs:533-546) with `range` instead
 * of 16 and `step` instead of 2:
 *   x_lo = po.x + clamp(-range_x + cmv.col/8, mvx_min/8, mvx_max/8)   etc.
 * with (mvx_min..mvy_max) = get_mv_range(w_in_b, h_in_b, po/4, blk_w, blk_h),
 * starting from (best_mv = 0, lowest_cost = MAX).
 * Results: out[job*R1C_SB_RESULTS + r1c_sb_result_offset(bsize) + k], k =
 * raster index of the block inside the region; sizes not in bsize_mask are
 * left untouched.
 * Precondition on cmv: the search centre must be a legal MV for every block of the region, i.e.
 * inside get_mv_range of each of them (true for the centres the encoder uses: 0, or a coarse MV
 * that full_search itself produced for the enclosing block).  r1c_me_sb_batch checks it
 * (R1C_ERR_ARG); the _dev forms cannot look at device job lists and do NOT: for a centre whose
 * whole window cmv +- range lies outside a block's range the reference clamps x_lo == x_hi to
 * the boundary and still evaluates that column, whereas these kernels mask every candidate and
 * return (mv 0, cost UINT64_MAX) for that block.  Callers with such centres clamp them first
 * (r1c_mv_range) or use r1c_full_search_batch, which reproduces the clamped window. */
#define R1C_SB_RESULTS 849
typedef struct {
  int32_t po_x, po_y;   /* region origin, multiple of 4, region inside the frame */
  uint32_t bsize_mask;  /* bit b = emit BlockSize b (sizes up to 64x64 only) */
  r1c_mv cmv;           /* search centre */
  r1c_mv pmv[2];
} r1c_sb_job;
typedef struct {
  int32_t w_in_b, h_in_b; /* frame size in 4x4 units (fi.w_in_b / fi.h_in_b) */
  int32_t range_x, range_y;
  int32_t step;           /* 1 or 2 */
  uint32_t lambda;
  int32_t allow_high_precision_mv;
} r1c_sb_params;
int r1c_sb_result_offset(int bsize); /* -1 for sizes above 64x64 */
int r1c_sb_result_count(int bsize);
int r1c_me_sb_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                    const r1c_sb_params *params, const r1c_sb_job *jobs, int64_t n,
                    int bit_depth, r1c_me_result *results);
int r1c_me_sb_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                        const r1c_sb_params *params, const r1c_sb_job *d_jobs, int64_t n,
                        int bit_depth, r1c_me_result *d_results);
/* Allocation-free variant of r1c_me_sb_batch_dev.  It has NO counterpart in the reference; it
 * was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs
 * and results as r1c_me_sb_batch_dev, but the search kernel's scratch (partial keys per
 * (superblock, chunk of candidate rows); org sums for u16 planes) lives in one file-scope
 * `static __device__` array of R1C_SB_NOALLOC_SCRATCH_BYTES (8 MiB, held by every GPU that loads
 * the library until exit) instead of the context's grow-on-demand scratch, and the call makes no
 * cudaMalloc / cudaFree / allocating library call.  Limits: only what the second-generation
 * kernel takes (step 1, lambda <= 60000, or <= 16000 on 12-bit planes; anything else
 * -> R1C_ERR_UNIMPLEMENTED, there is no fallback kernel here) and
 * n * chunks * 849 * 8 (+ n * 849 * 4 for u16 planes up to 10 bits) <= R1C_SB_NOALLOC_SCRATCH_BYTES -- e.g. all
 * 510 superblocks of a 1080p 8-bit frame at +-64 (2 chunks, 6.9 MB); more -> R1C_ERR_NOMEM with
 * the sizes in r1c_last_error().  All calls on a device share the array: calls that overlap on
 * the GPU (different streams or contexts) race; calls on one stream are ordered.  Nothing is
 * carried from one call to the next (every slot read is written by the same call first). */
#define R1C_SB_NOALLOC_SCRATCH_BYTES (8u << 20)
int r1c_me_sb_batch_dev_noalloc(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                const r1c_sb_params *params, const r1c_sb_job *d_jobs, int64_t n,
                                int bit_depth, r1c_me_result *d_results);

/* The full-pel stages of FullSearch motion estimation on the frame pyramid, batched over
 * blocks (window arithmetic of the reference on the host, searches = r1c_full_search_batch):
 *   ss4: estimate_motion_ss4, src/me.rs:1023-1075 -- quarter-res planes, window
 *        +-192*me_range_scale x +-64*me_range_scale luma px clipped by get_mv_range, >> 2;
 *        lambda = (me_lambda*256/16*0.125) as u32; result = best_mv * 4
 *   ss2: estimate_motion_ss2 + FullSearch::me_ss2, :280-333, 597-650 -- half-res planes,
 *        +-16 px around each of <= 3 predictor MVs (pmv_valid[i*3+k] != 0), best carried
 *        across the rounds; lambda = (me_lambda*256/4*0.125) as u32; result = best_mv * 2
 *   full: estimate_motion + FullSearch::full_pixel_me, :335-385, 522-565 -- +-16 px, step 2,
 *        around each of <= 7 candidate MVs; lambda = (me_lambda*256*0.5) as u32
 * jobs[i] = the block's offset in 4x4 units BEFORE adjust_bo (:993-1003) and its BlockSize;
 * a block without any valid predictor keeps MotionVector::default(). */
typedef struct {
  int32_t bo_x, bo_y;
  int32_t bsize;
} r1c_block_job;
int r1c_estimate_motion_ss4_batch(r1c_ctx *ctx, const r1c_plane *org_qres, const r1c_plane *ref_qres,
                                  int w_in_b, int h_in_b, const r1c_block_job *jobs, int64_t n,
                                  int me_range_scale, double me_lambda, int allow_high_precision_mv,
                                  int bit_depth, r1c_mv *out);
int r1c_estimate_motion_ss2_batch(r1c_ctx *ctx, const r1c_plane *org_hres, const r1c_plane *ref_hres,
                                  int w_in_b, int h_in_b, const r1c_block_job *jobs, int64_t n,
                                  const r1c_mv *pmvs, const uint8_t *pmv_valid, double me_lambda,
                                  int allow_high_precision_mv, int bit_depth, r1c_mv *out);
int r1c_estimate_motion_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int w_in_b,
                              int h_in_b, const r1c_block_job *jobs, int64_t n, const r1c_mv *cmvs,
                              const uint8_t *cmv_valid, double me_lambda, int allow_high_precision_mv,
                              int bit_depth, r1c_mv *out);

/* The motion-vector field: FrameMotionVectors (src/me.rs:32-62) = cols x rows r1c_mv in 4x4
 * units, row-major, owned by the caller.
 *   save_block_motion      src/encoder.rs:1432-1444: mvs[i] written to every cell block i covers
 *                          (jobs of one call must not overlap: the reference is sequential)
 *   get_subset_predictors  src/me.rs:82-174: per block the zero MV, its coarse MVs (cmvs[i*7 + k],
 *                          k < n_cmvs[i]) quantised to full-pel, EPZS subsets A/B from `field` and
 *                          C from `prev_field` (NULL = no reference frame); predictors[i*17 ..]
 *                          holds counts[i] entries in the reference's order, the rest repeats the
 *                          last one (the layout r1c_diamond_search_batch takes with n_pred = 17).
 * Jobs with bo outside [0, cols) x [0, rows) (or a bad bsize): R1C_ERR_ARG from the host forms;
 * the _dev forms skip them on the device (nothing written / counts[i] = 0, zero predictors). */
int r1c_save_block_motion_batch(r1c_ctx *ctx, r1c_mv *field, int cols, int rows,
                                const r1c_block_job *jobs, const r1c_mv *mvs, int64_t n);
int r1c_save_block_motion_batch_dev(r1c_ctx *ctx, r1c_mv *d_field, int cols, int rows,
                                    const r1c_block_job *d_jobs, const r1c_mv *d_mvs, int64_t n);
int r1c_subset_predictors_batch(r1c_ctx *ctx, const r1c_mv *field, const r1c_mv *prev_field, int cols,
                                int rows, const r1c_block_job *jobs, int64_t n, const r1c_mv *cmvs,
                                const uint8_t *n_cmvs, r1c_mv *predictors, int32_t *counts);
int r1c_subset_predictors_batch_dev(r1c_ctx *ctx, const r1c_mv *d_field, const r1c_mv *d_prev_field,
                                    int cols, int rows, const r1c_block_job *d_jobs, int64_t n,
                                    const r1c_mv *d_cmvs, const uint8_t *d_n_cmvs,
                                    r1c_mv *d_predictors, int32_t *d_counts);

/* Sub-pel candidate costs (get_mv_rd_cost with a temp plane, src/me.rs:787-838):
 * predict_inter (put_8tap, REGULAR/REGULAR = fi.default_filter) of the block at
 * each candidate MV, then SAD|SATD + rate.  n_cand candidates per job, laid out
 * cand[job*n_cand + k]; cost UINT64_MAX for MVs outside the job's range. */
typedef struct {
  int32_t po_x, po_y;
  int32_t bsize;
  uint32_t lambda;
  r1c_mv pmv[2];
  int32_t mvx_min, mvx_max, mvy_min, mvy_max;
  int32_t use_satd;
  int32_t allow_high_precision_mv;
} r1c_subpel_job;
int r1c_subpel_cost_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                          const r1c_subpel_job *jobs, int64_t n, const r1c_mv *cand, int n_cand,
                          int bit_depth, uint64_t *costs);
/* The sub-pel tail of MotionEstimation::motion_estimation for FullSearch
 * (src/me.rs:232-271): when use_satd, lowest_cost is first re-evaluated with
 * SATD at best_mv (full-pel path, :235-252); then telescopic_subpel_search
 * (:858-941): steps 8, 4, 2 (and 1 with high-precision MVs), 8 neighbours per
 * step in (i, j) raster order, strict `<`.  results[i] is in/out. */
int r1c_subpel_refine_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                            const r1c_subpel_job *jobs, int64_t n, int bit_depth,
                            r1c_me_result *results);
int r1c_subpel_refine_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                const r1c_subpel_job *d_jobs, int64_t n, int bit_depth,
                                r1c_me_result *d_results);
/* diamond_me_search (src/me.rs:693-785) with get_best_predictor (:655-691), the
 * search of DiamondSearch::{full_pixel_me, sub_pixel_me, me_ss2} (:391-520):
 * the best of `n_pred` predictor MVs per job (predictors[job*n_pred + k]; pad a
 * shorter list by repeating an entry -- ties keep the first) becomes the centre,
 * then 4-point diamonds of radius 16 -> 8 on the full-pel path (candidate block
 * at po + mv/8, truncating) or, with `subpixel`, 4 -> 2 (-> 1 with
 * allow_high_precision_mv) through predict_inter.  Pattern order (1,0),(0,1),
 * (-1,0),(0,-1) in (row, col), strict `<`, centre kept on ties.  The predictor
 * lists themselves (get_subset_predictors, :82-174) depend on the encoder's
 * neighbour state and are built by the caller.  results[i] is output only. */
int r1c_diamond_search_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                             const r1c_subpel_job *jobs, int64_t n, int subpixel,
                             const r1c_mv *predictors, int n_pred, int bit_depth,
                             r1c_me_result *results);
int r1c_diamond_search_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                 const r1c_subpel_job *d_jobs, int64_t n, int subpixel,
                                 int allow_high_precision_mv, const r1c_mv *d_predictors,
                                 int n_pred, int bit_depth, r1c_me_result *d_results);

/* ------------------------------------------------------------------------
 * Whole-frame motion estimation: MotionEstimation::motion_estimation for
 * FullSearch (src/me.rs:193-271) applied to EVERY block of every requested
 * BlockSize on its natural grid, one call per (frame, reference) pair --
 * the batched form of the loops in build_coarse_pmvs (src/encoder.rs:
 * 2727-2742) and of the per-block calls of the RDO loop.  Per block:
 *   full-pel : full_search over FullSearch::full_pixel_me's window
 *              (src/me.rs:533-546, cmv = 0, `range` instead of 16, step 1),
 *              SAD cost, from (best_mv = 0, lowest_cost = MAX);
 *   refine 1 : + the tail of motion_estimation (src/me.rs:232-271): SATD
 *              re-evaluation of the full-pel winner when use_satd, then
 *              telescopic_subpel_search (steps 8, 4, 2 (,1)).
 * pmv = {0, 0} for both predictors (what the lookahead passes).  This is full_search over
 * +-range at step 1 around cmv = 0 for every block, NOT a replay of the encoder's per-block
 * calls (FullSearch::full_pixel_me is +-16 at step 2 around a per-block cmv with the caller's
 * pmv, src/me.rs:533-565): those go through r1c_me_sb_batch (step, cmv, pmv per region) or
 * r1c_estimate_motion_batch (per block).
 * Result layout: for each BlockSize b of bsize_mask in enum order, the
 * (width / bw) x (height / bh) whole blocks in raster order:
 *   results[r1c_frame_me_offset(...b) + by * cols + bx].
 * ---------------------------------------------------------------------- */
typedef struct {
  int32_t width, height;    /* frame size in pixels, multiples of 4 */
  int32_t range_x, range_y; /* full-pel search range in pixels */
  uint32_t lambda;
  uint32_t bsize_mask;      /* sizes up to 64x64 only */
  int32_t refine;           /* 0: full-pel only, 1: + SATD / sub-pel tail */
  int32_t use_satd;         /* fi.config.speed_settings.use_satd_subpel */
  int32_t allow_high_precision_mv;
  /* multi-GPU sharding by superblock row: only blocks whose
   * origin lies in superblock rows [sb_row_lo, sb_row_hi) are computed; the
   * other entries of `results` are left untouched.  0, 0 = the whole frame. */
  int32_t sb_row_lo, sb_row_hi;
  /* ... and in superblock columns [sb_col_lo, sb_col_hi): with the rows, an AV1 tile (a rectangle of
   * superblocks, src/tiling/tiler.rs:87-109).  0, 0 = all columns. */
  int32_t sb_col_lo, sb_col_hi;
} r1c_frame_me_params;
/* number of result entries of the whole frame; offset / grid of one size
 * (returns -1 for sizes outside the mask) */
int64_t r1c_frame_me_count(const r1c_frame_me_params *p);
int64_t r1c_frame_me_offset(const r1c_frame_me_params *p, int bsize, int32_t *cols, int32_t *rows);
/* `results` is host memory; when it is page-locked (cudaHostAlloc / cudaHostRegister) the
 * device writes into it directly, otherwise through the library's pinned stage + a memcpy. */
int r1c_frame_me(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                 const r1c_frame_me_params *p, int bit_depth, r1c_me_result *results);
int r1c_frame_me_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                     const r1c_frame_me_params *p, int bit_depth, r1c_me_result *d_results);
/* The results of the tile `p` describes, packed: d_compact[i] = d_results[slot_i], slots in the
 * order BlockSize, block row, block column (the order of the frame layout restricted to the
 * tile); r1c_frame_me_shard_count() entries.  One contiguous device-to-host copy then returns a
 * rank's share of a frame (the tile's slots are not contiguous in the frame layout). */
int64_t r1c_frame_me_shard_count(r1c_ctx *ctx, const r1c_frame_me_params *p);
int r1c_frame_me_compact_dev(r1c_ctx *ctx, const r1c_frame_me_params *p,
                             const r1c_me_result *d_results, r1c_me_result *d_compact);
/* A lookahead window in one call: pairs[i] = one (frame, reference) pair, all with the same
 * geometry and parameters -- what the lookahead of src/api/internal.rs:557-573 computes frame by
 * frame.  The pairs go to TWO streams in turn (the context's own and an internal second one with
 * its own scratch, created on first use), so that the tail of one frame's kernels is filled by
 * the next frame's and, with host buffers, the copies of one frame overlap the kernels of
 * another: +13 % over calling r1c_frame_me(_dev) in a loop on one B200.
 *   r1c_frame_me_window_dev  planes resident, pairs[i].results = DEVICE array (frame layout);
 *                            only enqueues; work queued later on the context stream waits for
 *                            both streams
 *   r1c_frame_me_window      pairs[i].results = HOST array (page-locked: DMA target, else staged);
 *                            org_host / ref_host, when not NULL, are uploaded into the planes
 *                            first (whole padded plane, as r1c_plane_upload_async); blocking
 * Distinct pairs must not share a plane that is uploaded by the call. */
typedef struct {
  const r1c_plane *org, *ref;
  const void *org_host, *ref_host; /* r1c_frame_me_window only; NULL = the plane is resident */
  void *results;                   /* r1c_me_result array: device (_dev) or host */
} r1c_frame_pair;
int r1c_frame_me_window(r1c_ctx *ctx, const r1c_frame_pair *pairs, int64_t n,
                        const r1c_frame_me_params *p, int bit_depth);
int r1c_frame_me_window_dev(r1c_ctx *ctx, const r1c_frame_pair *pairs, int64_t n,
                            const r1c_frame_me_params *p, int bit_depth);
/* plane upload without the trailing synchronisation (pinned or pageable host
 * memory; ordered on the context stream) */
int r1c_plane_upload_async(r1c_ctx *ctx, r1c_plane *plane, const void *host_data);
/* rows [row_lo, row_hi) of the ALLOCATION only (row 0 = the first padding row; host_data still
 * points at row 0 of the whole padded plane).  A rank that searches one tile of a frame needs
 * the tile's rows of `org` and those rows +- (range_y + 16) of `ref`
 * (rav1e_b200.shard.tile_plane_rows), not the whole planes. */
int r1c_plane_upload_rows_async(r1c_ctx *ctx, r1c_plane *plane, const void *host_data,
                                int64_t row_lo, int64_t row_hi);
/* the same for a rectangle of the allocation: rows [row_lo, row_hi) x columns [col_lo, col_hi)
 * (column 0 = the first padding column; columns in pixels, <= stride).  What a rank needs of a
 * frame for one tile of a tile GRID: the tile +- (range + 16) in both directions
 * (rav1e_b200.shard.tile_plane_rows / tile_plane_cols).  One strided DMA (cudaMemcpy2DAsync). */
int r1c_plane_upload_rect_async(r1c_ctx *ctx, r1c_plane *plane, const void *host_data,
                                int64_t row_lo, int64_t row_hi, int64_t col_lo, int64_t col_hi);

/* ------------------------------------------------------------------------
 * Block importances of the lookahead:
 * ContextInner::compute_block_importances, src/api/internal.rs:823-1075.
 *   r1c_importance_propagate  one (frame, unique reference) pair of the loop :897-1047.
 *       frame / ref            luma planes of the input frame and of the reference frame
 *       mvs                    fi.lookahead_mvs[mv_index]: MI-granular field, mv_cols x mv_rows,
 *                              row-major (FrameMotionVectors, src/me.rs:32-62); block (x, y) reads
 *                              mvs[(2 y) * mv_cols + 2 x]
 *       intra_costs            fi.lookahead_intra_costs (r1c_lookahead_intra_costs), w*h u32
 *       importances            fi.block_importances of THIS frame (what the future propagated)
 *       n_refs                 unique_indices.len()
 *       ref_importances        block_importances of the reference frame, in/out; the shares are
 *                              added in the reference's order (raster over source blocks; TL, TR,
 *                              BL, BR), so the f32 sums are the reference's bit for bit
 *     An MV that moves its 8x8 block out of the allocated reference plane is R1C_ERR_BOUNDS
 *     (the reference's get_satd asserts); the _dev form clamps such blocks for memory safety.
 *   r1c_importance_finalize   :1054-1071, importance = log2(1 + importance / intra_cost) or 0;
 *     log2 is computed in f64 and rounded once; the reference takes log2f from the libm it is
 *     linked with (not correctly rounded, platform-dependent): the two agree to 1 ulp of f32
 *     (tested against glibc; 90+ % of the cells are equal).
 * ---------------------------------------------------------------------- */
int r1c_importance_propagate(r1c_ctx *ctx, const r1c_plane *frame, const r1c_plane *ref,
                             const r1c_mv *mvs, int mv_cols, int mv_rows,
                             const uint32_t *intra_costs, const float *importances, int n_refs,
                             int w_in_imp_b, int h_in_imp_b, int bit_depth, float *ref_importances);
int r1c_importance_propagate_dev(r1c_ctx *ctx, const r1c_plane *frame, const r1c_plane *ref,
                                 const r1c_mv *d_mvs, int mv_cols, int mv_rows,
                                 const uint32_t *d_intra_costs, const float *d_importances,
                                 int n_refs, int w_in_imp_b, int h_in_imp_b, int bit_depth,
                                 float *d_ref_importances);
int r1c_importance_finalize(r1c_ctx *ctx, const uint32_t *intra_costs, float *importances, int64_t n);
int r1c_importance_finalize_dev(r1c_ctx *ctx, const uint32_t *d_intra_costs, float *d_importances,
                                int64_t n);

#ifdef __cplusplus
}
#endif
#endif /* RAV1E_CUDA_H */
