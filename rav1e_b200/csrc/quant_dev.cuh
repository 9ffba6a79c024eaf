// quant_dev.cuh -- the quantizer of one warp's blocks held in shared memory (QuantizationContext::quantize,
// reference src/quantize.rs:257-316; divu_pair :140-153; dequantize :318-334), shared by quantize_kernel (quant.cu)
// and the fused encode_tx_block kernel (enc_tx.cuh).  See quant.cu for the lane / level_mode scheme.
#pragma once

#include "common.cuh"

namespace quant {

struct QuantConsts {
  r1c_quant_ctx qc;
  const uint16_t *scan;
  int32_t deadzone;  // (ac_quant - ac_offset_eob) rounded up and shifted by log_tx_scale, src/quantize.rs:265-267
  int nonneg;        // 0 <= ac_offset0 <= ac_offset1 < ac_offset0 + ac_quant: the magnitude fast path applies
};

static inline QuantConsts make_consts(const r1c_quant_ctx &qc, const uint16_t *d_scan) {
  QuantConsts q;
  q.qc = qc;
  q.scan = d_scan;
  // (ac_quant as usize - ac_offset_eob as usize).align_power_of_two_and_shift(log_tx_scale), src/quantize.rs:265-267
  q.deadzone = (int32_t)(((uint64_t)qc.ac_quant - (uint64_t)qc.ac_offset_eob + (1ull << qc.log_tx_scale) - 1) >> qc.log_tx_scale);
  q.nonneg = qc.ac_offset0 >= 0 && qc.ac_offset1 >= qc.ac_offset0 && (uint32_t)qc.ac_offset1 < qc.ac_quant &&
             qc.ac_quant < (1u << 30) && qc.dc_offset >= 0 && (uint32_t)qc.dc_offset < qc.dc_quant && qc.dc_quant < (1u << 30);
  return q;
}

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

// The same walk for quantization contexts as QuantizationContext::update leaves them (a.nonneg: 0 <= ac_offset0 <=
// ac_offset1 < ac_quant, 0 <= dc_offset < dc_quant), as straight-line code: magnitudes are divided once (divu_pair is an
// exact division, the reference's test_divu_pair; the quotient with the larger offset is the one with the smaller
// offset plus a carry out of its remainder), `level0 > 0 / > 1` are comparisons with d / 2 d, a negative coefficient
// takes ac_offset0 under both modes, the DC (:278-280) is the same formula with its own divisor and one offset, and a
// coefficient past the end of block is one whose quotients are zeroed.  With these offsets a lane's map
// {mode before} -> {mode after} is the identity, "set" (f0 = f1 = 1) or "reset" (f0 = f1 = 0), never a flip
// (q1 >= q0), so a lane's incoming mode is the kind of the closest earlier set / reset lane of its block.
template <int LPB, int NCH, int LOGW>
__device__ __forceinline__ unsigned long long quantize_lanes_fast(int32_t *sb, int l, int sh, const QuantConsts &a, bool want_dist) {
  constexpr uint32_t BMASK = LPB == 32 ? 0xffffffffu : ((1u << LPB) - 1u);
  const int lts = (int)a.qc.log_tx_scale;
  auto sidx = [](int p) { return LOGW < 0 ? p : p + (p >> (LOGW < 0 ? 0 : LOGW)); };
  constexpr bool ROLLED = NCH >= 16;  // see quantize_lanes
  constexpr int UNR = ROLLED ? 4 : NCH;
  int32_t raw_keep[ROLLED ? 1 : NCH];
  int pos_keep[ROLLED ? 1 : NCH];
  int eob = 1;
#pragma unroll UNR
  for (int c = 0; c < NCH; c++) {
    const int k = c * LPB + l;
    const int p = sidx((int)__ldg(a.scan + k));
    const int32_t r = sb[p];
    if (!ROLLED) pos_keep[c] = p, raw_keep[c] = r;
    if (k >= 1 && (r < 0 ? -r : r) >= a.deadzone) eob = k;
  }
#pragma unroll
  for (int m = LPB >> 1; m; m >>= 1) eob = max(eob, __shfl_xor_sync(0xffffffffu, eob, m, LPB));

  const uint32_t ma = a.qc.ac_mul_add[0], mb = a.qc.ac_mul_add[1], ms = a.qc.ac_mul_add[2];
  const uint32_t d_ac = a.qc.ac_quant, off_ac = (uint32_t)a.qc.ac_offset0, doff_ac = (uint32_t)(a.qc.ac_offset1 - a.qc.ac_offset0);
  const uint32_t blk_all = BMASK << sh;               // this block's lanes in a ballot
  const uint32_t blk_below = ((1u << l) - 1u) << sh;  // the earlier ones
  const uint32_t dq_mask = (1u << lts) - 1u;
  // chunks that lie past the end of block of EVERY block of the warp are zeroed without going through the
  // quantizer (residuals of a real encode leave most of a large block's scan behind the end of block)
  int weob = eob;
#pragma unroll
  for (int m = 16; m >= LPB; m >>= 1) weob = max(weob, __shfl_xor_sync(0xffffffffu, weob, m));
  __syncwarp();
  uint32_t mode = 1;  // level_mode at the start of the chunk (the same in all lanes of a block)
  unsigned long long dist = 0;
  // rolled loops stop at the warp's last live chunk (the tail is zeroed below); unrolled ones test every chunk
  const int nlive = ROLLED ? min(NCH, weob / LPB + 1) : NCH;
#pragma unroll UNR
  for (int c = 0; c < nlive; c++) {
    const int k = c * LPB + l;
    const int pos_c = ROLLED ? sidx((int)__ldg(a.scan + k)) : pos_keep[c];
    const int32_t raw_c = ROLLED ? sb[pos_c] : raw_keep[c];  // each position is read and rewritten by its one lane
    if (!ROLLED && c * LPB > weob) {  // warp-uniform
      sb[pos_c] = 0;
      if (want_dist) dist += (unsigned long long)(long long)(int32_t)((uint32_t)raw_c * (uint32_t)raw_c);
      continue;
    }
    const int32_t coeff = (int32_t)((uint32_t)raw_c << lts);
    const bool neg = coeff < 0;
    const uint32_t y0 = (uint32_t)(neg ? -coeff : coeff);
    uint32_t d = d_ac, off0 = off_ac, doff = doff_ac, m0 = ma, m1 = mb, m2 = ms;
    const bool dc = c == 0 && l == 0;
    if (c == 0) {  // lane 0 of the first chunk swaps its constants
      d = dc ? a.qc.dc_quant : d, off0 = dc ? (uint32_t)a.qc.dc_offset : off0, doff = dc ? 0u : doff;
      m0 = dc ? a.qc.dc_mul_add[0] : m0, m1 = dc ? a.qc.dc_mul_add[1] : m1, m2 = dc ? a.qc.dc_mul_add[2] : m2;
    }
    const uint32_t t0 = y0 + off0;
    uint32_t ua = udivu(t0, m0, m1, m2);                      // (|coeff| + offset0) / d; 0 when coeff == 0 (offset0 < d)
    uint32_t ub = ua + (t0 - ua * d + doff >= d ? 1u : 0u);   // (|coeff| + offset1) / d
    const bool in = k <= eob;                                 // src/quantize.rs:313-315 zeroes the rest
    ua = in ? ua : 0u;
    ub = (in && !neg) ? ub : ua;
    const uint32_t q0 = y0 >= 2 * d ? ub : ua;                // `level0 > 1 - level_mode`, mode 0 / 1
    const uint32_t q1 = y0 >= d ? ub : ua;
    // mode 0: `else if qcoeff > 1 { level_mode = 1 }`; mode 1: `qcoeff == 0` -> 0; the DC and the tail leave it alone
    const bool f0 = !neg && q0 > 1 && !dc;
    const bool f1 = q1 != 0 || !in || dc;
    const uint32_t A = __ballot_sync(0xffffffffu, f0), B = __ballot_sync(0xffffffffu, f1);
    const uint32_t fixed = ~(A ^ B) & blk_all;  // lanes that set (A = B = 1) or reset (A = B = 0) the mode
    const uint32_t below = fixed & blk_below;
    const uint32_t m_in = below ? (A >> (31 - __clz(below))) & 1u : mode;
    mode = fixed ? (A >> (31 - __clz(fixed))) & 1u : mode;
    const uint32_t qm = m_in ? q1 : q0;
    const int32_t q = neg ? -(int32_t)qm : (int32_t)qm;
    sb[pos_c] = q;
    if (want_dist) {
      const int32_t r = (int32_t)((uint32_t)q * (pos_c == 0 ? a.qc.dc_quant : d_ac) + ((uint32_t)(q >> 31) & dq_mask)) >> lts;
      const uint32_t e = (uint32_t)raw_c - (uint32_t)r;
      dist += (unsigned long long)(long long)(int32_t)(e * e);  // (c * c) as u64, wrapping i32 product
    }
  }
  if (ROLLED) {
    for (int c = nlive; c < NCH; c++) {
      const int pos_c = sidx((int)__ldg(a.scan + c * LPB + l));
      const uint32_t raw_c = (uint32_t)sb[pos_c];
      sb[pos_c] = 0;
      if (want_dist) dist += (unsigned long long)(long long)(int32_t)(raw_c * raw_c);
    }
  }
  return dist;
}

// Quantizes in place the block at `sb` (row-major coefficients; LOGW < 0: dense, else rows of 2^LOGW coefficients
// W + 1 words apart, the layout the 2-D transforms leave) with the LPB lanes l = 0 .. LPB-1 that share it; `sh` is the
// warp lane of the block's lane 0.  ALL 32 lanes of the warp must call it (ballots).  Returns this lane's share of
// sum((coeff - dequantized)^2) when want_dist (else 0); the caller adds the LPB shares.
template <int LPB, int NCH, int LOGW>
__device__ __forceinline__ unsigned long long quantize_lanes(int32_t *sb, int l, int sh, const QuantConsts &a, bool want_dist) {
  if (a.nonneg) return quantize_lanes_fast<LPB, NCH, LOGW>(sb, l, sh, a, want_dist);
  constexpr uint32_t BMASK = LPB == 32 ? 0xffffffffu : ((1u << LPB) - 1u);
  const int lts = (int)a.qc.log_tx_scale;
  auto sidx = [](int p) { return LOGW < 0 ? p : p + (p >> (LOGW < 0 ? 0 : LOGW)); };
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
    const int p = sidx((int)__ldg(a.scan + k));
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
    const int pos_c = ROLLED ? sidx((int)__ldg(a.scan + k)) : pos_keep[c];
    const int32_t raw_c = ROLLED ? sb[pos_c] : raw_keep[c];  // each position is read and rewritten by its one lane
    const int32_t coeff = (int32_t)((uint32_t)raw_c << lts);
    // this lane's result under level_mode 0 / 1 and its transition map f0 / f1 = next mode from mode 0 / 1;
    // the DC and the zeroed tail leave the mode alone (f0 = 0, f1 = 1)
    int32_t q0, q1;
    bool f0 = false, f1 = true;
    if (k == 0) {
      // DC, src/quantize.rs:278-280
      q0 = q1 = divu_pair(coeff + signum(coeff) * a.qc.dc_offset, a.qc.dc_mul_add[0], a.qc.dc_mul_add[1], a.qc.dc_mul_add[2]);
    } else if (k <= eob) {
      // src/quantize.rs:293-311, any offsets (contexts that did not come from update())
      const int32_t level0 = divu_pair(coeff, ma, mb, ms);
      const int32_t sg = signum(coeff);
      const int32_t qa = divu_pair(coeff + sg * a.qc.ac_offset0, ma, mb, ms);
      const int32_t qb = divu_pair(coeff + sg * a.qc.ac_offset1, ma, mb, ms);
      q0 = level0 > 1 ? qb : qa;  // `level0 > 1 - level_mode`
      q1 = level0 > 0 ? qb : qa;
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
    if (want_dist) {
      const int32_t r = dequant1(q, pos_c == 0 ? a.qc.dc_quant : a.qc.ac_quant, lts);
      const uint32_t d = (uint32_t)raw_c - (uint32_t)r;
      dist += (unsigned long long)(long long)(int32_t)(d * d);  // (c * c) as u64, wrapping i32 product
    }
  }
  return dist;
}

}  // namespace quant
