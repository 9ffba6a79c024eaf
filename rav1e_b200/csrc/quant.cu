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
#include "common.cuh"

namespace {

struct QuantArgs {
  r1c_quant_ctx qc;
  const uint16_t *scan;
  const int32_t *coeffs;
  int32_t *qcoeffs, *rcoeffs;
  uint64_t *tx_dist;
  long long n;
  int area, coded, dist_bits;
  int32_t deadzone;
  int nonneg;  // 0 <= ac_offset0 <= ac_offset1 < ac_offset0 + ac_quant: the magnitude fast path applies
};

// divu_pair, src/quantize.rs:140-153: ((a * |x| + b) >> 32) >> shift with the sign of x
__device__ __forceinline__ int32_t divu_pair(int32_t x, uint32_t a, uint32_t b, uint32_t shift) {
  const uint32_t y = (uint32_t)(x < 0 ? -x : x);
  const int32_t q = (int32_t)((uint32_t)(((unsigned long long)a * y + b) >> 32) >> shift);
  return x < 0 ? -q : q;
}
// the magnitude part of divu_pair
__device__ __forceinline__ uint32_t udivu(uint32_t y, uint32_t a, uint32_t b, uint32_t shift) {
  return (uint32_t)(((unsigned long long)a * y + b) >> 32) >> shift;
}
__device__ __forceinline__ int32_t signum(int32_t v) { return (v > 0) - (v < 0); }
// dequantize, src/quantize.rs:329-332 (wrapping i32 product, arithmetic shift)
__device__ __forceinline__ int32_t dequant1(int32_t c, uint32_t quant, int lts) {
  return (int32_t)((uint32_t)c * quant + (uint32_t)((c >> 31) & ((1 << lts) - 1))) >> lts;
}

// LPB lanes per block, NCH = coded / LPB chunks of scan positions; a lane keeps its NCH coefficients in registers
// between the end-of-block search and the quantizer pass
template <int LPB, int NCH>
__global__ void __launch_bounds__(256) quantize_kernel(const __grid_constant__ QuantArgs a) {
  constexpr int BPW = 32 / LPB;  // blocks per warp
  const int lane = threadIdx.x & 31, l = lane & (LPB - 1), sh = lane & ~(LPB - 1);
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long b = warp * BPW + lane / LPB;
  const bool live = b < a.n;  // lanes of a block that does not exist go through the motions on zeros
  const int lts = (int)a.qc.log_tx_scale;
  constexpr uint32_t BMASK = LPB == 32 ? 0xffffffffu : ((1u << LPB) - 1u);
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

  // ---- eob: last scan index k >= 1 with |coeff| >= deadzone, else 1 (src/quantize.rs:268-276) ----
  // up to 8 chunks per lane: both loops fully unrolled, the lane's coefficients and scan positions stay in registers
  // between them.  16 and 32 chunks (512 / 1024 coded coefficients): rolled loops that read them again from shared
  // memory / the scan table -- the unrolled form needs 64 - 124 registers and does not fit the instruction cache
  constexpr bool ROLLED = NCH >= 16;
  constexpr int UNR = ROLLED ? 4 : NCH;
  int32_t raw_keep[ROLLED ? 1 : NCH];
  int pos_keep[ROLLED ? 1 : NCH];
  int eob = 1;
#pragma unroll UNR
  for (int c = 0; c < NCH; c++) {
    const int k = c * LPB + l;
    const int p = (int)__ldg(a.scan + k);
    const int32_t r = sb[p];
    if (!ROLLED) pos_keep[c] = p, raw_keep[c] = r;
    if (k >= 1 && (r < 0 ? -r : r) >= a.deadzone) eob = k;
  }
#pragma unroll
  for (int m = LPB >> 1; m; m >>= 1) eob = max(eob, __shfl_xor_sync(0xffffffffu, eob, m, LPB));

  const uint32_t ma = a.qc.ac_mul_add[0], mb = a.qc.ac_mul_add[1], ms = a.qc.ac_mul_add[2];
  __syncwarp();
  uint32_t mode = 1;  // level_mode at the start of the chunk
  unsigned long long dist = 0;
#pragma unroll UNR
  for (int c = 0; c < NCH; c++) {
    const int k = c * LPB + l;
    const int pos_c = ROLLED ? (int)__ldg(a.scan + k) : pos_keep[c];
    const int32_t raw_c = ROLLED ? sb[pos_c] : raw_keep[c];  // each position is read and rewritten by its one lane
    const int32_t coeff = (int32_t)((uint32_t)raw_c << lts);
    // this lane's result under level_mode 0 / 1 and its transition map f0 / f1 = next mode from mode 0 / 1;
    // the DC and the zeroed tail leave the mode alone (f0 = 0, f1 = 1)
    int32_t q0, q1;
    bool f0 = false, f1 = true;
    if (k == 0 && !a.nonneg) {
      // DC, src/quantize.rs:278-280
      q0 = q1 = divu_pair(coeff + signum(coeff) * a.qc.dc_offset, a.qc.dc_mul_add[0], a.qc.dc_mul_add[1], a.qc.dc_mul_add[2]);
    } else if (k <= eob) {
      // src/quantize.rs:293-311
      if (a.nonneg) {
        // offsets >= 0 (always, after update()): divide magnitudes, apply the sign once.  A negative coefficient has
        // level0 <= 0, so it takes ac_offset0 under both modes and can only reset the mode (the tests are signed)
        // divu_pair is an exact division (Robison's multiply-add form; the reference's test_divu_pair), so
        // level0 > 0 / > 1 are comparisons with d / 2 d, and the quotient with the larger offset is the one with the
        // smaller offset plus a carry out of its remainder (0 <= offset1 - offset0 < d): one wide multiply per coefficient.
        // The DC (:278-280) is the same formula with its own divisor and one offset for both roles, so lane 0 of
        // chunk 0 only swaps its constants instead of sending the warp through a second division
        const bool dc = c == 0 && k == 0;
        const uint32_t d = dc ? a.qc.dc_quant : a.qc.ac_quant;
        const uint32_t off0 = (uint32_t)(dc ? a.qc.dc_offset : a.qc.ac_offset0);
        const uint32_t doff = dc ? 0u : (uint32_t)(a.qc.ac_offset1 - a.qc.ac_offset0);
        const uint32_t y0 = (uint32_t)(coeff < 0 ? -coeff : coeff);
        const uint32_t t0 = y0 + off0;
        uint32_t ua = c == 0 ? udivu(t0, dc ? a.qc.dc_mul_add[0] : ma, dc ? a.qc.dc_mul_add[1] : mb, dc ? a.qc.dc_mul_add[2] : ms)
                             : udivu(t0, ma, mb, ms);
        uint32_t ub = ua + (t0 - ua * d + doff >= d ? 1u : 0u);
        ua = y0 ? ua : 0u;
        ub = y0 ? ub : 0u;
        const bool neg = coeff < 0;
        q0 = neg ? -(int32_t)ua : (int32_t)(y0 >= 2 * d ? ub : ua);
        q1 = neg ? -(int32_t)ua : (int32_t)(y0 >= d ? ub : ua);
      } else {
        const int32_t level0 = divu_pair(coeff, ma, mb, ms);
        const int32_t sg = signum(coeff);
        const int32_t qa = divu_pair(coeff + sg * a.qc.ac_offset0, ma, mb, ms);
        const int32_t qb = divu_pair(coeff + sg * a.qc.ac_offset1, ma, mb, ms);
        q0 = level0 > 1 ? qb : qa;  // `level0 > 1 - level_mode`
        q1 = level0 > 0 ? qb : qa;
      }
      // mode 0: `else if qcoeff > 1 { level_mode = 1 }`; mode 1: `qcoeff == 0` -> 0, otherwise it stays 1;
      // the DC is outside the state machine
      f0 = k != 0 && q0 > 1;
      f1 = k == 0 || q1 != 0;
    } else {
      q0 = q1 = 0;  // src/quantize.rs:313-315
    }
    // maps of the block's lanes as two bit masks; a lane's incoming mode = the mode left by the closest earlier
    // lane whose map is constant (f0 == f1), else the mode carried into the chunk
    const uint32_t A = (__ballot_sync(0xffffffffu, f0) >> sh) & BMASK, B = (__ballot_sync(0xffffffffu, f1) >> sh) & BMASK;
    uint32_t m_in;
    if (__any_sync(0xffffffffu, f0 && !f1)) {
      // a map that flips the mode (q0 > 1 but q1 == 0: not seen with offset1 >= offset0, handled all the same):
      // compose the 2-bit maps with an inclusive shuffle scan
      uint32_t inc = (f0 ? 1u : 0u) | (f1 ? 2u : 0u);
#pragma unroll
      for (int d = 1; d < LPB; d <<= 1) {
        const uint32_t prev = __shfl_up_sync(0xffffffffu, inc, d, LPB);
        if (l >= d) inc = ((inc >> (prev & 1u)) & 1u) | (((inc >> ((prev >> 1) & 1u)) & 1u) << 1);
      }
      uint32_t exc = __shfl_up_sync(0xffffffffu, inc, 1, LPB);
      if (l == 0) exc = 2u;
      m_in = (exc >> mode) & 1u;
      mode = (__shfl_sync(0xffffffffu, inc, LPB - 1, LPB) >> mode) & 1u;
    } else {
      const uint32_t fixed = ~(A ^ B) & BMASK;  // lanes that set (A = B = 1) or reset (A = B = 0) the mode
      const uint32_t below = fixed & ((1u << l) - 1u);
      m_in = below ? (A >> (31 - __clz(below))) & 1u : mode;
      mode = fixed ? (A >> (31 - __clz(fixed))) & 1u : mode;
    }
    const int32_t q = m_in ? q1 : q0;
    sb[pos_c] = q;
    if (a.tx_dist) {
      const int32_t r = dequant1(q, pos_c == 0 ? a.qc.dc_quant : a.qc.ac_quant, lts);
      const uint32_t d = (uint32_t)raw_c - (uint32_t)r;
      dist += (unsigned long long)(long long)(int32_t)(d * d);  // (c * c) as u64, wrapping i32 product
    }
  }
  __syncwarp();
#pragma unroll
  for (int i = 0; i < BPW * CODED / 32; i++) {
    const int e = i * 32 + lane, blk = e / CODED, off = e - blk * CODED;
    const long long gb = warp * BPW + blk;
    if (gb < a.n) {
      const int32_t q = sw[e];
      a.qcoeffs[gb * CODED + off] = q;
      if (a.rcoeffs) a.rcoeffs[gb * CODED + off] = dequant1(q, off == 0 ? a.qc.dc_quant : a.qc.ac_quant, lts);
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
  a.qc = *qc;
  a.scan = d_scan;
  a.coeffs = d_coeffs;
  a.qcoeffs = d_qcoeffs;
  a.rcoeffs = d_rcoeffs;
  a.tx_dist = d_tx_dist;
  a.n = n;
  a.area = r1c_tx_width(tx_size) * r1c_tx_height(tx_size);
  a.coded = coded_of(tx_size);
  a.dist_bits = 2 * (3 - (int)qc->log_tx_scale);
  // (ac_quant as usize - ac_offset_eob as usize).align_power_of_two_and_shift(log_tx_scale), src/quantize.rs:265-267
  a.deadzone = (int32_t)(((uint64_t)qc->ac_quant - (uint64_t)qc->ac_offset_eob + (1ull << qc->log_tx_scale) - 1) >> qc->log_tx_scale);
  a.nonneg = qc->ac_offset0 >= 0 && qc->ac_offset1 >= qc->ac_offset0 && (uint32_t)(qc->ac_offset1 - qc->ac_offset0) < qc->ac_quant &&
             qc->ac_quant < (1u << 30) && qc->dc_offset >= 0 && (uint32_t)qc->dc_offset < qc->dc_quant && qc->dc_quant < (1u << 30);
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
