// tx.cu -- C-ABI entry points of the forward transforms (forward_transform,
// reference src/transform/mod.rs:556-566 -> FwdTxfm2D::fht forward.rs:1804-1899)
// and dispatch to the per-(size, type) kernels instantiated in tx_inst_*.cu.
#include "tx2d.cuh"

namespace tx {
#define TX_DECL(NAME) extern const fht_launch_fn NAME[9];
TX_DECL(k_tx_4x4) TX_DECL(k_tx_8x8) TX_DECL(k_tx_16x16) TX_DECL(k_tx_32x32) TX_DECL(k_tx_64x64)
TX_DECL(k_tx_4x8) TX_DECL(k_tx_8x4) TX_DECL(k_tx_8x16) TX_DECL(k_tx_16x8) TX_DECL(k_tx_16x32)
TX_DECL(k_tx_32x16) TX_DECL(k_tx_32x64) TX_DECL(k_tx_64x32) TX_DECL(k_tx_4x16) TX_DECL(k_tx_16x4)
TX_DECL(k_tx_8x32) TX_DECL(k_tx_32x8) TX_DECL(k_tx_16x64) TX_DECL(k_tx_64x16)

// order of TxSize, src/transform/mod.rs:225-247
static const fht_launch_fn *const SIZE_TABLES[R1C_TX_SIZES_ALL] = {
    k_tx_4x4,  k_tx_8x8,   k_tx_16x16, k_tx_32x32, k_tx_64x64, k_tx_4x8,  k_tx_8x4,
    k_tx_8x16, k_tx_16x8,  k_tx_16x32, k_tx_32x16, k_tx_32x64, k_tx_64x32, k_tx_4x16,
    k_tx_16x4, k_tx_8x32,  k_tx_32x8,  k_tx_16x64, k_tx_64x16};

// (column 1-D, row 1-D) of each TxType, src/transform/mod.rs:196-213; 3 = FLIPADST
static const unsigned char COL_T[R1C_TX_TYPES] = {0, 1, 0, 1, 3, 0, 3, 1, 3, 2, 0, 2, 1, 2, 3, 2};
static const unsigned char ROW_T[R1C_TX_TYPES] = {0, 0, 1, 1, 0, 3, 3, 3, 1, 2, 2, 0, 2, 1, 2, 3};

// FWD_SHIFT_*, src/transform/forward.rs:22-40; rows = (bit_depth - 8) / 2
static const signed char SH_A[3][3] = {{3, 0, 0}, {2, 0, 1}, {0, 0, 3}};
static const signed char SH_B[3][3] = {{4, -1, 0}, {2, 0, 1}, {0, 0, 3}};
static const signed char SH_C[3][3] = {{4, -2, 0}, {2, 0, 0}, {0, 0, 2}};
static const signed char SH_D[3][3] = {{4, -1, -2}, {2, 0, -1}, {0, 0, 1}};
static const signed char (*shift_rows(int tx_size))[3] {
  switch (tx_size) {
    case R1C_TX_4X4: return SH_A;
    case R1C_TX_32X32: case R1C_TX_16X32: case R1C_TX_32X16: case R1C_TX_16X64: case R1C_TX_64X16: return SH_C;
    case R1C_TX_64X64: case R1C_TX_32X64: case R1C_TX_64X32: return SH_D;
    default: return SH_B;
  }
}

static fht_launch_fn lookup(int tx_size, int tx_type, int *col_flip, int *row_flip) {
  if (tx_size < 0 || tx_size >= R1C_TX_SIZES_ALL || tx_type < 0 || tx_type >= R1C_TX_TYPES) return nullptr;
  int ct = COL_T[tx_type], rt = ROW_T[tx_type];
  *col_flip = ct == 3;
  *row_flip = rt == 3;
  if (ct == 3) ct = TX1_ADST;
  if (rt == 3) rt = TX1_ADST;
  return SIZE_TABLES[tx_size][ct * 3 + rt];
}

static int launch(r1c_ctx *ctx, int tx_size, int tx_type, int bit_depth, FhtArgs &a, int fused_pb) {
  R1C_REQUIRE(ctx, bit_depth == 8 || bit_depth == 10 || bit_depth == 12, R1C_ERR_ARG, "bit_depth %d", bit_depth);
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL && tx_type >= 0 && tx_type < R1C_TX_TYPES,
              R1C_ERR_ARG, "bad (tx_size, tx_type) = (%d, %d)", tx_size, tx_type);
  fht_launch_fn fn = lookup(tx_size, tx_type, &a.col_flip, &a.row_flip);
  R1C_REQUIRE(ctx, fn != nullptr, R1C_ERR_UNIMPLEMENTED,
              "forward transform (size %d, type %d) is unimplemented in the reference", tx_size, tx_type);
  const signed char *sh = shift_rows(tx_size)[(bit_depth - 8) / 2];
  a.sh0 = sh[0];
  a.sh1 = sh[1];
  a.sh2 = sh[2];
  // every table has entry 0 >= 0 (left shift) and entry 1 <= 0 (rounding right shift)
  a.ls0 = sh[0];
  a.rs1 = -sh[1];
  a.rnd1 = (1 << a.rs1) >> 1;
  a.ls2 = sh[2] > 0 ? sh[2] : 0;
  a.rs2 = sh[2] < 0 ? -sh[2] : 0;
  a.rnd2 = (1 << a.rs2) >> 1;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  r1c_prof_begin(ctx, R1C_PROF_TX);
  const int rc = fn(ctx, a, fused_pb);
  r1c_prof_end(ctx, R1C_PROF_TX);
  return rc;
}
}  // namespace tx

void r1c_fwd_shift_row(int tx_size, int bit_depth, int out[3]) {
  const signed char *sh = tx::shift_rows(tx_size)[(bit_depth - 8) / 2];
  out[0] = sh[0], out[1] = sh[1], out[2] = sh[2];
}

extern "C" int r1c_fwd_txfm_valid(int tx_size, int tx_type) {
  int cf, rf;
  return tx::lookup(tx_size, tx_type, &cf, &rf) != nullptr;
}

extern "C" int r1c_fwd_txfm_batch_dev(r1c_ctx *ctx, int tx_size, int tx_type, int bit_depth,
                                      const int16_t *d_residual, int32_t *d_coeffs, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_residual && d_coeffs)), R1C_ERR_ARG, "null residual/coeff array");
  tx::FhtArgs a = {};
  a.residual = d_residual;
  a.coeffs = d_coeffs;
  a.n = n;
  return tx::launch(ctx, tx_size, tx_type, bit_depth, a, 0);
}

extern "C" int r1c_fwd_txfm_batch(r1c_ctx *ctx, int tx_size, int tx_type, int bit_depth,
                                  const int16_t *residual, int32_t *coeffs, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (residual && coeffs)), R1C_ERR_ARG, "null residual/coeff array");
  R1C_REQUIRE(ctx, r1c_tx_width(tx_size) > 0, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, r1c_fwd_txfm_valid(tx_size, tx_type), tx_type >= 0 && tx_type < R1C_TX_TYPES ? R1C_ERR_UNIMPLEMENTED : R1C_ERR_ARG,
              "forward transform (size %d, type %d) not available", tx_size, tx_type);
  if (n == 0) return R1C_OK;
  const size_t area = (size_t)r1c_tx_width(tx_size) * r1c_tx_height(tx_size);
  const size_t ib = (size_t)n * area * 2, ob = (size_t)n * area * 4;
  const void *h_in;
  void *h_out, *d_in, *d_out;
  int rc, staged;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_host_src(ctx, 0, residual, ib, &h_in))) return rc;  // page-locked buffers are the DMA source / target themselves
  if ((rc = r1c_host_dst(ctx, 1, coeffs, ob, &h_out, &staged))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, ib, &d_in))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_in, h_in, ib, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_fwd_txfm_batch_dev(ctx, tx_size, tx_type, bit_depth, (const int16_t *)d_in, (int32_t *)d_out, n))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (staged) memcpy(coeffs, h_out, ob);
  return R1C_OK;
}

extern "C" int r1c_forward_transform(r1c_ctx *ctx, const int16_t *residual, int32_t *coeffs,
                                     int tx_size, int tx_type, int bit_depth) {
  return r1c_fwd_txfm_batch(ctx, tx_size, tx_type, bit_depth, residual, coeffs, 1);
}

static int check_tx_planes(r1c_ctx *ctx, const r1c_plane *input, const r1c_plane *pred, int bit_depth) {
  if (!ctx || !input || !pred) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, input->px_bytes == pred->px_bytes, R1C_ERR_ARG, "input/pred pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, input->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, input->px_bytes);
  return R1C_OK;
}

extern "C" int r1c_diff_fwd_txfm_batch_dev(r1c_ctx *ctx, const r1c_plane *input, const r1c_plane *pred,
                                           const r1c_tx_job *d_jobs, int64_t n, int tx_size, int tx_type,
                                           int bit_depth, int32_t *d_coeffs) {
  int rc = check_tx_planes(ctx, input, pred, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_coeffs)), R1C_ERR_ARG, "null job/coeff array");
  tx::FhtArgs a = {};
  a.coeffs = d_coeffs;
  a.n = n;
  a.input = make_view(input);
  a.pred = make_view(pred);
  a.jobs = d_jobs;
  return tx::launch(ctx, tx_size, tx_type, bit_depth, a, input->px_bytes);
}

extern "C" int r1c_diff_fwd_txfm_batch(r1c_ctx *ctx, const r1c_plane *input, const r1c_plane *pred,
                                       const r1c_tx_job *jobs, int64_t n, int tx_size, int tx_type,
                                       int bit_depth, int32_t *coeffs) {
  int rc = check_tx_planes(ctx, input, pred, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (jobs && coeffs)), R1C_ERR_ARG, "null job/coeff array");
  R1C_REQUIRE(ctx, r1c_tx_width(tx_size) > 0, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, r1c_fwd_txfm_valid(tx_size, tx_type), tx_type >= 0 && tx_type < R1C_TX_TYPES ? R1C_ERR_UNIMPLEMENTED : R1C_ERR_ARG,
              "forward transform (size %d, type %d) not available", tx_size, tx_type);
  if (n == 0) return R1C_OK;
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size);
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, region_in_plane(input->cfg, jobs[i].in_x, jobs[i].in_y, w, h) && region_in_plane(pred->cfg, jobs[i].pred_x, jobs[i].pred_y, w, h),
                R1C_ERR_BOUNDS, "tx job %lld leaves the allocated plane", (long long)i);
  const size_t jb = (size_t)n * sizeof(r1c_tx_job), ob = (size_t)n * w * h * 4;
  void *h_jobs, *h_out, *d_jobs, *d_out;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, ob, &h_out))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  memcpy(h_jobs, jobs, jb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_diff_fwd_txfm_batch_dev(ctx, input, pred, (const r1c_tx_job *)d_jobs, n, tx_size, tx_type, bit_depth, (int32_t *)d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(coeffs, h_out, ob);
  return R1C_OK;
}
