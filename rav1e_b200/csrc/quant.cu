// quant.cu -- batched quantize / dequantize with the tx-domain distortion fused in
// (reference src/quantize.rs:123-153 divu_gen / divu_pair, :212-255 QuantizationContext::update,
// :257-316 QuantizationContext::quantize, :318-334 dequantize; src/encoder.rs:1210-1224 tx_dist).
//
// quantize walks the coefficients in scan order with a two-state rounding machine (`level_mode`,
// src/quantize.rs:287-311): the offset of coefficient k depends on the mode left by coefficient
// k-1.  Here LPB lanes (16 for 4x4 blocks, else 32) take LPB consecutive scan positions at once:
// a lane computes its coefficient's result under BOTH modes, which also gives the lane's
// transition map {0,1} -> {0,1} (2 bits); two ballots collect the maps and a lane's incoming mode
// is the mode left by the closest earlier lane whose map is constant (a shuffle scan composes the
// maps in the general, mode-flipping case).  Blocks are independent: a warp takes 32 / LPB blocks, a CTA 8 warps.
// The end-of-block search (rposition over |coeff| >= deadzone) is a max-reduction over the
// same lanes; a lane keeps its coefficients in registers for the second pass.
// Bound: HBM, 4 B read + 4 B written per coefficient (+4 with rcoeffs); `scan` stays in L1.
#include "quant_dev.cuh"

namespace {
using quant::dequant1;

struct QuantArgs {
  quant::QuantConsts q;
  const int32_t *coeffs;
  int32_t *qcoeffs, *rcoeffs;
  uint64_t *tx_dist;
  long long n;
  int area, coded, dist_bits;
};

// LPB lanes per block, NCH = coded / LPB chunks of scan positions; a lane keeps its NCH coefficients in registers
// between the end-of-block search and the quantizer pass
template <int LPB, int NCH>
__global__ void __launch_bounds__(256) quantize_kernel(const __grid_constant__ QuantArgs a) {
  constexpr int BPW = 32 / LPB;  // blocks per warp
  const int lane = threadIdx.x & 31, l = lane & (LPB - 1), sh = lane & ~(LPB - 1);
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long b = warp * BPW + lane / LPB;
  const bool live = b < a.n;  // lanes of a block that does not exist go through the motions on zeros
  const int lts = (int)a.q.qc.log_tx_scale;
  // The warp's blocks go through shared memory: coalesced in, gathered / scattered by scan position there (a scan
  // sends neighbouring lanes to different rows of the block: one L1 wavefront per lane when done on global memory),
  // coalesced out.  Quantized values overwrite the coefficients in place.
  constexpr int CODED = LPB * NCH;
  __shared__ int32_t s_buf[8][BPW * CODED];
  int32_t *sw = s_buf[threadIdx.x >> 5], *sb = sw + (lane / LPB) * CODED;
#pragma unroll
  for (int i = 0; i < BPW * CODED / 32; i++) {
    const int e = i * 32 + lane, blk = e / CODED, off = e - blk * CODED;
    const long long gb = warp * BPW + blk;
    sw[e] = gb < a.n ? __ldg(a.coeffs + gb * a.area + off) : 0;
  }
  __syncwarp();

  const unsigned long long part = quant::quantize_lanes<LPB, NCH, -1>(sb, l, sh, a.q, a.tx_dist != nullptr);
  unsigned long long dist = part;
  __syncwarp();
#pragma unroll
  for (int i = 0; i < BPW * CODED / 32; i++) {
    const int e = i * 32 + lane, blk = e / CODED, off = e - blk * CODED;
    const long long gb = warp * BPW + blk;
    if (gb < a.n) {
      const int32_t q = sw[e];
      a.qcoeffs[gb * CODED + off] = q;
      if (a.rcoeffs) a.rcoeffs[gb * CODED + off] = dequant1(q, off == 0 ? a.q.qc.dc_quant : a.q.qc.ac_quant, lts);
    }
  }
  if (a.tx_dist) {
#pragma unroll
    for (int m = LPB >> 1; m; m >>= 1) dist += __shfl_xor_sync(0xffffffffu, dist, m, LPB);
    if (l == 0 && live) a.tx_dist[b] = (dist + (1ull << (a.dist_bits - 1))) >> a.dist_bits;
  }
}

__global__ void dequantize_kernel(const int32_t *q, int32_t *r, long long total, int coded, uint32_t dc_quant,
                                  uint32_t ac_quant, int lts) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const bool dc = (i % coded) == 0;
  r[i] = dequant1(__ldg(q + i), dc ? dc_quant : ac_quant, lts);
}

inline int coded_of(int tx_size) {
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size);
  return (w < 32 ? w : 32) * (h < 32 ? h : 32);  // av1_get_coded_tx_size, src/encoder.rs:1143
}

// divu_gen, src/quantize.rs:123-138
void divu_gen(uint32_t d, uint32_t out[3]) {
  uint32_t m = 31;
  while (!((d >> m) & 1u)) m--;
  if ((d & (d - 1)) == 0) {
    out[0] = out[1] = 0xFFFFFFFFu;
  } else {
    const unsigned long long t = (1ull << (m + 32)) / d;
    const unsigned long long r = (t * d + d) & 0xFFFFFFFFull;
    if (r <= (1ull << m)) {
      out[0] = (uint32_t)t + 1;
      out[1] = 0;
    } else {
      out[0] = out[1] = (uint32_t)t;
    }
  }
  out[2] = m;
}

}  // namespace

extern "C" int r1c_log_tx_scale(int tx_size) {
  if (tx_size < 0 || tx_size >= R1C_TX_SIZES_ALL) return -1;
  const int area = r1c_tx_width(tx_size) * r1c_tx_height(tx_size);
  return (area > 256) + (area > 1024);
}

extern "C" int r1c_quant_ctx_update(r1c_quant_ctx *qc, int tx_size, int is_intra, uint32_t dc_quant, uint32_t ac_quant) {
  if (!qc || tx_size < 0 || tx_size >= R1C_TX_SIZES_ALL || dc_quant == 0 || ac_quant == 0 || dc_quant > 32767 ||
      ac_quant > 32767)  // dc_q / ac_q return i16 (src/quantize.rs:41,52); 0 would be lossless (unsupported there too)
    return R1C_ERR_ARG;
  qc->log_tx_scale = (uint32_t)r1c_log_tx_scale(tx_size);
  qc->dc_quant = dc_quant;
  divu_gen(dc_quant, qc->dc_mul_add);
  qc->ac_quant = ac_quant;
  divu_gen(ac_quant, qc->ac_mul_add);
  qc->dc_offset = (int32_t)dc_quant * (is_intra ? 109 : 108) / 256;
  qc->ac_offset0 = (int32_t)ac_quant * (is_intra ? 98 : 97) / 256;
  qc->ac_offset1 = (int32_t)ac_quant * (is_intra ? 109 : 108) / 256;
  qc->ac_offset_eob = (int32_t)ac_quant * (is_intra ? 88 : 44) / 256;
  return R1C_OK;
}

extern "C" int r1c_quantize_batch_dev(r1c_ctx *ctx, const r1c_quant_ctx *qc, int tx_size, const uint16_t *d_scan,
                                      const int32_t *d_coeffs, int32_t *d_qcoeffs, int32_t *d_rcoeffs,
                                      uint64_t *d_tx_dist, int64_t n) {
  if (!ctx || !qc) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_scan && d_coeffs && d_qcoeffs)), R1C_ERR_ARG, "null scan / coefficient array");
  R1C_REQUIRE(ctx, qc->ac_quant > 0 && qc->dc_quant > 0 && (int)qc->log_tx_scale == r1c_log_tx_scale(tx_size), R1C_ERR_ARG,
              "quantization context was not updated for tx_size %d", tx_size);
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  QuantArgs a;
  a.q = quant::make_consts(*qc, d_scan);
  a.coeffs = d_coeffs;
  a.qcoeffs = d_qcoeffs;
  a.rcoeffs = d_rcoeffs;
  a.tx_dist = d_tx_dist;
  a.n = n;
  a.area = r1c_tx_width(tx_size) * r1c_tx_height(tx_size);
  a.coded = coded_of(tx_size);
  a.dist_bits = 2 * (3 - (int)qc->log_tx_scale);
  r1c_prof_begin(ctx, R1C_PROF_QUANT);
  const unsigned g8 = (unsigned)((n + 7) / 8);
  switch (a.coded) {  // 4x4 .. 32x32 coded areas
    // small blocks share a warp (4 or 8 lanes each): the per-block steps (end-of-block reduction, ballots,
    // staging index arithmetic) are then spread over four to eight coefficients per lane instead of one or two
    case 16: quantize_kernel<4, 4><<<(unsigned)((n + 63) / 64), 256, 0, ctx->stream>>>(a); break;
    case 32: quantize_kernel<8, 4><<<(unsigned)((n + 31) / 32), 256, 0, ctx->stream>>>(a); break;
    case 64: quantize_kernel<8, 8><<<(unsigned)((n + 31) / 32), 256, 0, ctx->stream>>>(a); break;
    case 128: quantize_kernel<32, 4><<<g8, 256, 0, ctx->stream>>>(a); break;
    case 256: quantize_kernel<32, 8><<<g8, 256, 0, ctx->stream>>>(a); break;
    case 512: quantize_kernel<32, 16><<<g8, 256, 0, ctx->stream>>>(a); break;
    default: quantize_kernel<32, 32><<<g8, 256, 0, ctx->stream>>>(a); break;
  }
  r1c_prof_end(ctx, R1C_PROF_QUANT);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_dequantize_batch_dev(r1c_ctx *ctx, int tx_size, uint32_t dc_quant, uint32_t ac_quant,
                                        const int32_t *d_qcoeffs, int32_t *d_rcoeffs, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_qcoeffs && d_rcoeffs)), R1C_ERR_ARG, "null coefficient array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const int coded = coded_of(tx_size);
  const long long total = (long long)n * coded;
  r1c_prof_begin(ctx, R1C_PROF_QUANT);
  dequantize_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(d_qcoeffs, d_rcoeffs, total, coded, dc_quant,
                                                                              ac_quant, r1c_log_tx_scale(tx_size));
  r1c_prof_end(ctx, R1C_PROF_QUANT);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_quantize_batch(r1c_ctx *ctx, const r1c_quant_ctx *qc, int tx_size, const uint16_t *scan,
                                  const int32_t *coeffs, int32_t *qcoeffs, int64_t n) {
  if (!ctx || !qc) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (scan && coeffs && qcoeffs)), R1C_ERR_ARG, "null scan / coefficient array");
  if (n == 0) return R1C_OK;
  const int coded = coded_of(tx_size), area = r1c_tx_width(tx_size) * r1c_tx_height(tx_size);
  {  // scan[0] == 0 (the DC has its own quantizer) and every index is hit once
    unsigned char seen[1024] = {0};
    R1C_REQUIRE(ctx, scan[0] == 0, R1C_ERR_ARG, "scan[0] must be the DC coefficient");
    for (int k = 0; k < coded; k++) {
      R1C_REQUIRE(ctx, scan[k] < coded && !seen[scan[k]], R1C_ERR_ARG, "scan is not a permutation of 0..%d", coded - 1);
      seen[scan[k]] = 1;
    }
  }
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t cb = (size_t)n * area * 4, qb = (size_t)n * coded * 4, sb = (size_t)coded * 2;
  void *d_c, *d_q, *d_s, *h_q;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 0, cb, &d_c))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, qb, &d_q))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, sb, &d_s))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, qb, &h_q))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_c, coeffs, cb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_s, scan, sb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_quantize_batch_dev(ctx, qc, tx_size, (const uint16_t *)d_s, (const int32_t *)d_c, (int32_t *)d_q, nullptr,
                                   nullptr, n)))
    return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_q, d_q, qb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(qcoeffs, h_q, qb);
  return R1C_OK;
}

extern "C" int r1c_dequantize_batch(r1c_ctx *ctx, int tx_size, uint32_t dc_quant, uint32_t ac_quant,
                                    const int32_t *qcoeffs, int32_t *rcoeffs, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (qcoeffs && rcoeffs)), R1C_ERR_ARG, "null coefficient array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t qb = (size_t)n * coded_of(tx_size) * 4;
  void *d_q, *d_r, *h_r;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 0, qb, &d_q))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, qb, &d_r))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, qb, &h_r))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_q, qcoeffs, qb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_dequantize_batch_dev(ctx, tx_size, dc_quant, ac_quant, (const int32_t *)d_q, (int32_t *)d_r, n))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_r, d_r, qb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(rcoeffs, h_r, qb);
  return R1C_OK;
}
