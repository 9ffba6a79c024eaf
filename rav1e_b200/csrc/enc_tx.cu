// enc_tx.cu -- C-ABI entry points of the fused residual chain of encode_tx_block (reference
// src/encoder.rs:1143-1224) and dispatch to the per-(size, type) kernels instantiated in enc_tx_inst_*.cu.
#include "enc_tx.cuh"

namespace enc {
#define ENC_DECL(NAME) extern const enc_launch_fn NAME[9];
ENC_DECL(k_enc_4x4) ENC_DECL(k_enc_8x8) ENC_DECL(k_enc_16x16) ENC_DECL(k_enc_32x32)
ENC_DECL(k_enc_4x8) ENC_DECL(k_enc_8x4) ENC_DECL(k_enc_8x16) ENC_DECL(k_enc_16x8) ENC_DECL(k_enc_16x32)
ENC_DECL(k_enc_32x16) ENC_DECL(k_enc_4x16) ENC_DECL(k_enc_16x4) ENC_DECL(k_enc_8x32) ENC_DECL(k_enc_32x8)
ENC_DECL(k_enc_64x64) ENC_DECL(k_enc_32x64) ENC_DECL(k_enc_64x32) ENC_DECL(k_enc_16x64) ENC_DECL(k_enc_64x16)

// order of TxSize, src/transform/mod.rs:225-247; the five sizes with a 64-point side have their own kernel
// (enc_tx64.cuh: their forward transform leaves 64-wide rows while the quantizer and the inverse read the first
// min(W,32) * min(H,32) words, the inverse as 32-wide rows); their 64-point side is a DCT
static const enc_launch_fn *const SIZE_TABLES[R1C_TX_SIZES_ALL] = {
    k_enc_4x4,  k_enc_8x8,  k_enc_16x16, k_enc_32x32, k_enc_64x64, k_enc_4x8,   k_enc_8x4,
    k_enc_8x16, k_enc_16x8, k_enc_16x32, k_enc_32x16, k_enc_32x64, k_enc_64x32, k_enc_4x16,
    k_enc_16x4, k_enc_8x32, k_enc_32x8,  k_enc_16x64, k_enc_64x16};

// (column 1-D, row 1-D) of each TxType, src/transform/mod.rs:196-213; 3 = FLIPADST
static const unsigned char COL_T[R1C_TX_TYPES] = {0, 1, 0, 1, 3, 0, 3, 1, 3, 2, 0, 2, 1, 2, 3, 2};
static const unsigned char ROW_T[R1C_TX_TYPES] = {0, 0, 1, 1, 0, 3, 3, 3, 1, 2, 2, 0, 2, 1, 2, 3};

static enc_launch_fn lookup(int tx_size, int tx_type, int *col_flip, int *row_flip) {
  if (tx_size < 0 || tx_size >= R1C_TX_SIZES_ALL || tx_type < 0 || tx_type >= R1C_TX_TYPES) return nullptr;
  if (!SIZE_TABLES[tx_size]) return nullptr;
  int ct = COL_T[tx_type], rt = ROW_T[tx_type];
  *col_flip = ct == 3;
  *row_flip = rt == 3;
  if (ct == 3) ct = 1;
  if (rt == 3) rt = 1;
  return SIZE_TABLES[tx_size][ct * 3 + rt];
}

static itx::Range range_of(int bits) {  // clamp_value, src/transform/mod.rs:492-496
  itx::Range r;
  r.hi = (int)(((long long)1 << (bits - 1)) - 1);
  r.lo = (int)(-((long long)1 << (bits - 1)));
  return r;
}
}  // namespace enc

extern "C" int r1c_encode_tx_block_valid(int tx_size, int tx_type) {
  int cf, rf;
  return enc::lookup(tx_size, tx_type, &cf, &rf) != nullptr;
}

extern "C" int r1c_encode_tx_block_batch_dev(r1c_ctx *ctx, const r1c_plane *input, r1c_plane *rec, const r1c_tx_job *d_jobs,
                                             int64_t n, int tx_size, int tx_type, int bit_depth, const r1c_quant_ctx *qc,
                                             const uint16_t *d_scan, int reconstruct, int32_t *d_qcoeffs, uint64_t *d_tx_dist) {
  if (!ctx || !input || !rec || !qc) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, input->px_bytes == rec->px_bytes, R1C_ERR_ARG, "input/rec pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, input->px_bytes) && (bit_depth == 8 || bit_depth == 10 || bit_depth == 12), R1C_ERR_ARG,
              "bit_depth %d vs %d-byte pixels", bit_depth, input->px_bytes);
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL && tx_type >= 0 && tx_type < R1C_TX_TYPES, R1C_ERR_ARG,
              "bad (tx_size, tx_type) = (%d, %d)", tx_size, tx_type);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_scan && d_qcoeffs)), R1C_ERR_ARG, "null job / scan / qcoeff array");
  R1C_REQUIRE(ctx, qc->ac_quant > 0 && qc->dc_quant > 0 && (int)qc->log_tx_scale == r1c_log_tx_scale(tx_size), R1C_ERR_ARG,
              "quantization context was not updated for tx_size %d", tx_size);
  enc::EncArgs a = {};
  enc::enc_launch_fn fn = enc::lookup(tx_size, tx_type, &a.col_flip, &a.row_flip);
  R1C_REQUIRE(ctx, fn != nullptr, R1C_ERR_UNIMPLEMENTED,
              "encode_tx_block (size %d, type %d): a type / size pair the reference leaves unimplemented", tx_size, tx_type);
  if (n == 0) return R1C_OK;
  a.input = make_view(input);
  a.rec = make_view(rec);
  a.jobs = d_jobs;
  a.n = n;
  a.qcoeffs = d_qcoeffs;
  a.tx_dist = d_tx_dist;
  a.reconstruct = reconstruct != 0;
  a.px_bytes = input->px_bytes;
  int sh[3];
  r1c_fwd_shift_row(tx_size, bit_depth, sh);
  // every table has entry 0 >= 0 (left shift) and entry 1 <= 0 (rounding right shift), see tx.cu
  a.ls0 = sh[0];
  a.rs1 = -sh[1];
  a.rnd1 = (1 << a.rs1) >> 1;
  a.ls2 = sh[2] > 0 ? sh[2] : 0;
  a.rs2 = sh[2] < 0 ? -sh[2] : 0;
  a.rnd2 = (1 << a.rs2) >> 1;
  a.q = quant::make_consts(*qc, d_scan);
  a.dist_bits = 2 * (3 - (int)qc->log_tx_scale);
  a.rg_row = enc::range_of(bit_depth + 8);
  a.rg_col = enc::range_of(bit_depth + 6 > 16 ? bit_depth + 6 : 16);
  a.ish = r1c_inv_intermediate_shift(tx_size);
  a.irnd = (1 << a.ish) >> 1;
  a.maxv = (1 << bit_depth) - 1;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  r1c_prof_begin(ctx, R1C_PROF_TX);
  const int rc = fn(ctx, a);
  r1c_prof_end(ctx, R1C_PROF_TX);
  return rc;
}

extern "C" int r1c_encode_tx_block_batch(r1c_ctx *ctx, const r1c_plane *input, r1c_plane *rec, const r1c_tx_job *jobs, int64_t n,
                                         int tx_size, int tx_type, int bit_depth, const r1c_quant_ctx *qc, const uint16_t *scan,
                                         int reconstruct, int32_t *qcoeffs, uint64_t *tx_dist) {
  if (!ctx || !input || !rec || !qc) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (jobs && scan && qcoeffs)), R1C_ERR_ARG, "null job / scan / qcoeff array");
  R1C_REQUIRE(ctx, r1c_tx_width(tx_size) > 0, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, r1c_encode_tx_block_valid(tx_size, tx_type), tx_type >= 0 && tx_type < R1C_TX_TYPES ? R1C_ERR_UNIMPLEMENTED : R1C_ERR_ARG,
              "encode_tx_block (size %d, type %d): a type / size pair the reference leaves unimplemented", tx_size, tx_type);
  if (n == 0) return R1C_OK;
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size);
  const int coded = (w < 32 ? w : 32) * (h < 32 ? h : 32);  // av1_get_coded_tx_size, src/context.rs:1949-1956
  {  // scan[0] == 0 (the DC has its own quantizer) and every index is hit once
    unsigned char seen[1024] = {0};
    R1C_REQUIRE(ctx, scan[0] == 0, R1C_ERR_ARG, "scan[0] must be the DC coefficient");
    for (int k = 0; k < coded; k++) {
      R1C_REQUIRE(ctx, scan[k] < coded && !seen[scan[k]], R1C_ERR_ARG, "scan is not a permutation of 0..%d", coded - 1);
      seen[scan[k]] = 1;
    }
  }
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, region_in_plane(input->cfg, jobs[i].in_x, jobs[i].in_y, w, h) && region_in_plane(rec->cfg, jobs[i].pred_x, jobs[i].pred_y, w, h),
                R1C_ERR_BOUNDS, "tx job %lld leaves the allocated plane", (long long)i);
  const size_t jb = (size_t)n * sizeof(r1c_tx_job), qb = (size_t)n * coded * 4, sb = (size_t)coded * 2, db = (size_t)n * 8;
  void *h_jobs, *h_q, *h_d = nullptr, *d_jobs, *d_q, *d_s, *d_d = nullptr;
  int rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, qb, &h_q))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, qb, &d_q))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, sb, &d_s))) return rc;
  if (tx_dist) {
    if ((rc = r1c_stage_host(ctx, 2, db, &h_d))) return rc;
    if ((rc = r1c_scratch_dev(ctx, 3, db, &d_d))) return rc;
  }
  memcpy(h_jobs, jobs, jb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_s, scan, sb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_encode_tx_block_batch_dev(ctx, input, rec, (const r1c_tx_job *)d_jobs, n, tx_size, tx_type, bit_depth, qc,
                                          (const uint16_t *)d_s, reconstruct, (int32_t *)d_q, (uint64_t *)d_d)))
    return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_q, d_q, qb, cudaMemcpyDeviceToHost, ctx->stream));
  if (tx_dist) R1C_CUDA(ctx, cudaMemcpyAsync(h_d, d_d, db, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(qcoeffs, h_q, qb);
  if (tx_dist) memcpy(tx_dist, h_d, db);
  return R1C_OK;
}
