// itx.cu -- C-ABI entry points of the inverse transforms (inverse_transform_add, reference
// src/transform/mod.rs:568-580 -> InvTxfm2D::inv_txfm2d_add, src/transform/inverse.rs:1939-2028) and
// dispatch to the per-(size, type) kernels instantiated in itx_inst_*.cu.
#include "itx2d.cuh"

namespace itx {
#define ITX_DECL(NAME) extern const itx_launch_fn NAME[9];
ITX_DECL(k_itx_4x4) ITX_DECL(k_itx_8x8) ITX_DECL(k_itx_16x16) ITX_DECL(k_itx_32x32) ITX_DECL(k_itx_64x64)
ITX_DECL(k_itx_4x8) ITX_DECL(k_itx_8x4) ITX_DECL(k_itx_8x16) ITX_DECL(k_itx_16x8) ITX_DECL(k_itx_16x32)
ITX_DECL(k_itx_32x16) ITX_DECL(k_itx_32x64) ITX_DECL(k_itx_64x32) ITX_DECL(k_itx_4x16) ITX_DECL(k_itx_16x4)
ITX_DECL(k_itx_8x32) ITX_DECL(k_itx_32x8) ITX_DECL(k_itx_16x64) ITX_DECL(k_itx_64x16)

// order of TxSize, src/transform/mod.rs:225-247
static const itx_launch_fn *const SIZE_TABLES[R1C_TX_SIZES_ALL] = {
    k_itx_4x4,  k_itx_8x8,  k_itx_16x16, k_itx_32x32, k_itx_64x64, k_itx_4x8,  k_itx_8x4,
    k_itx_8x16, k_itx_16x8, k_itx_16x32, k_itx_32x16, k_itx_32x64, k_itx_64x32, k_itx_4x16,
    k_itx_16x4, k_itx_8x32, k_itx_32x8,  k_itx_16x64, k_itx_64x16};

// (column 1-D, row 1-D) of each TxType, src/transform/mod.rs:196-213; 3 = FLIPADST
static const unsigned char COL_T[R1C_TX_TYPES] = {0, 1, 0, 1, 3, 0, 3, 1, 3, 2, 0, 2, 1, 2, 3, 2};
static const unsigned char ROW_T[R1C_TX_TYPES] = {0, 0, 1, 1, 0, 3, 3, 3, 1, 2, 2, 0, 2, 1, 2, 3};

static itx_launch_fn lookup(int tx_size, int tx_type, int *col_flip, int *row_flip) {
  if (tx_size < 0 || tx_size >= R1C_TX_SIZES_ALL || tx_type < 0 || tx_type >= R1C_TX_TYPES) return nullptr;
  int ct = COL_T[tx_type], rt = ROW_T[tx_type];
  *col_flip = ct == 3;
  *row_flip = rt == 3;
  if (ct == 3) ct = I1_ADST;
  if (rt == 3) rt = I1_ADST;
  return SIZE_TABLES[tx_size][ct * 3 + rt];
}

// InvBlock::INTERMEDIATE_SHIFT, inverse.rs:1656-1665
static int intermediate_shift(int tx_size) {
  switch (tx_size) {
    case R1C_TX_4X4: case R1C_TX_4X8: case R1C_TX_8X4: return 0;
    case R1C_TX_16X16: case R1C_TX_16X64: case R1C_TX_64X16: case R1C_TX_64X64:
    case R1C_TX_32X32: case R1C_TX_8X32: case R1C_TX_32X8: return 2;
    default: return 1;
  }
}

static Range range_of(int bits) {  // clamp_value, src/transform/mod.rs:492-496
  Range r;
  r.hi = (int)(((long long)1 << (bits - 1)) - 1);
  r.lo = (int)(-((long long)1 << (bits - 1)));
  return r;
}
}  // namespace itx

int r1c_inv_intermediate_shift(int tx_size) { return itx::intermediate_shift(tx_size); }

extern "C" int r1c_inv_txfm_valid(int tx_size, int tx_type) {
  int cf, rf;
  return itx::lookup(tx_size, tx_type, &cf, &rf) != nullptr;
}

// dq: nullptr, or {dc_quant, ac_quant} for the fused dequantize
static int launch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *d_jobs, int64_t n, int tx_size, int tx_type,
                      int bit_depth, const int32_t *d_coeffs, const uint32_t *dq) {
  if (!ctx || !dst) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, dst->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, dst->px_bytes);
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL && tx_type >= 0 && tx_type < R1C_TX_TYPES, R1C_ERR_ARG,
              "bad (tx_size, tx_type) = (%d, %d)", tx_size, tx_type);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_coeffs)), R1C_ERR_ARG, "null job/coeff array");
  R1C_REQUIRE(ctx, !dq || (dq[0] && dq[1]), R1C_ERR_ARG, "zero quantizer");
  itx::ItxArgs a = {};
  itx::itx_launch_fn fn = itx::lookup(tx_size, tx_type, &a.col_flip, &a.row_flip);
  R1C_REQUIRE(ctx, fn != nullptr, R1C_ERR_UNIMPLEMENTED, "inverse transform (size %d, type %d) is unimplemented in the reference",
              tx_size, tx_type);
  a.coeffs = d_coeffs;
  a.dst = make_view(dst);
  a.jobs = d_jobs;
  a.n = n;
  a.rg_row = itx::range_of(bit_depth + 8);
  a.rg_col = itx::range_of(bit_depth + 6 > 16 ? bit_depth + 6 : 16);
  a.sh = itx::intermediate_shift(tx_size);
  a.rnd = (1 << a.sh) >> 1;
  a.maxv = (1 << bit_depth) - 1;
  if (dq) {
    a.dq_on = 1;
    a.dq_dc = dq[0];
    a.dq_ac = dq[1];
    a.dq_lts = r1c_log_tx_scale(tx_size);
  }
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  r1c_prof_begin(ctx, R1C_PROF_ITX);
  const int rc = fn(ctx, a, dst->px_bytes);
  r1c_prof_end(ctx, R1C_PROF_ITX);
  return rc;
}

extern "C" int r1c_inv_txfm_add_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *d_jobs, int64_t n,
                                          int tx_size, int tx_type, int bit_depth, const int32_t *d_rcoeffs) {
  return launch_dev(ctx, dst, d_jobs, n, tx_size, tx_type, bit_depth, d_rcoeffs, nullptr);
}

extern "C" int r1c_dequant_inv_txfm_add_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *d_jobs, int64_t n,
                                                  int tx_size, int tx_type, int bit_depth, const int32_t *d_qcoeffs,
                                                  uint32_t dc_quant, uint32_t ac_quant) {
  const uint32_t dq[2] = {dc_quant, ac_quant};
  return launch_dev(ctx, dst, d_jobs, n, tx_size, tx_type, bit_depth, d_qcoeffs, dq);
}

static int launch_host(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *jobs, int64_t n, int tx_size, int tx_type,
                       int bit_depth, const int32_t *rcoeffs, const uint32_t *dq) {
  if (!ctx || !dst) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (jobs && rcoeffs)), R1C_ERR_ARG, "null job/coeff array");
  R1C_REQUIRE(ctx, r1c_tx_width(tx_size) > 0, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, r1c_inv_txfm_valid(tx_size, tx_type), tx_type >= 0 && tx_type < R1C_TX_TYPES ? R1C_ERR_UNIMPLEMENTED : R1C_ERR_ARG,
              "inverse transform (size %d, type %d) not available", tx_size, tx_type);
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, dst->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, dst->px_bytes);
  if (n == 0) return R1C_OK;
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size);
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, region_in_plane(dst->cfg, jobs[i].x, jobs[i].y, w, h), R1C_ERR_BOUNDS,
                "itx job %lld leaves the allocated plane", (long long)i);
  const size_t coded = (size_t)(w < 32 ? w : 32) * (h < 32 ? h : 32);
  const size_t jb = (size_t)n * sizeof(r1c_itx_job), cb = (size_t)n * coded * 4;
  void *h_jobs, *h_c, *d_jobs, *d_c;
  int rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, cb, &h_c))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, cb, &d_c))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_c, rcoeffs, cb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_c, h_c, cb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_dev(ctx, dst, (const r1c_itx_job *)d_jobs, n, tx_size, tx_type, bit_depth, (const int32_t *)d_c, dq))) return rc;
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_inv_txfm_add_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *jobs, int64_t n, int tx_size,
                                      int tx_type, int bit_depth, const int32_t *rcoeffs) {
  return launch_host(ctx, dst, jobs, n, tx_size, tx_type, bit_depth, rcoeffs, nullptr);
}

extern "C" int r1c_dequant_inv_txfm_add_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_itx_job *jobs, int64_t n, int tx_size,
                                              int tx_type, int bit_depth, const int32_t *qcoeffs, uint32_t dc_quant,
                                              uint32_t ac_quant) {
  const uint32_t dq[2] = {dc_quant, ac_quant};
  return launch_host(ctx, dst, jobs, n, tx_size, tx_type, bit_depth, qcoeffs, dq);
}
