/*
 * r1o_quant.c -- CPU ORACLE (test infrastructure only, see r1o.h): quantize / dequantize.
 *
 * Restates, statement by statement, the reference's
 *   divu_gen / divu_pair            src/quantize.rs:123-153
 *   QuantizationContext::update     src/quantize.rs:212-255 (from the dc_q / ac_q VALUES: the
 *                                   qindex lookup tables :318-... stay with the caller)
 *   QuantizationContext::quantize   src/quantize.rs:257-316
 *   dequantize                      src/quantize.rs:318-334
 * with the coefficient type of the encoder's call site (T = i32, src/encoder.rs:1144-1170).
 * Parity UNPINNED: the reference holds no known-answer test for quantize (only test_divu_pair,
 * src/quantize.rs:160-168, which tests/test_oracle_props.py repeats against this file).
 */
#include "r1o.h"

/* get_log_tx_scale, src/quantize.rs:34-39 */
int r1o_log_tx_scale(int tx_size) {
  const int area = r1o_tx_width(tx_size) * r1o_tx_height(tx_size);
  return (area > 256) + (area > 1024);
}

/* divu_gen, src/quantize.rs:123-138 */
void r1o_divu_gen(uint32_t d, uint32_t out[3]) {
  const uint64_t nbits = 32;
  const uint64_t m = nbits - (uint64_t)__builtin_clz(d) - 1;
  if ((d & (d - 1)) == 0) {
    out[0] = 0xFFFFFFFFu;
    out[1] = 0xFFFFFFFFu;
    out[2] = (uint32_t)m;
  } else {
    const uint64_t dd = d;
    const uint64_t t = (1ull << (m + nbits)) / dd;
    const uint64_t r = (t * dd + dd) & ((1ull << nbits) - 1);
    if (r <= (1ull << m)) {
      out[0] = (uint32_t)t + 1;
      out[1] = 0;
    } else {
      out[0] = (uint32_t)t;
      out[1] = (uint32_t)t;
    }
    out[2] = (uint32_t)m;
  }
}

/* divu_pair, src/quantize.rs:140-153 */
int32_t r1o_divu_pair(int32_t x, const uint32_t d[3]) {
  const uint64_t y = (uint64_t)(int64_t)(x < 0 ? -x : x);
  const uint64_t a = d[0], b = d[1], shift = d[2];
  const int32_t q = (int32_t)(((a * y + b) >> 32) >> shift);
  return x < 0 ? -q : q;
}

/* QuantizationContext::update, src/quantize.rs:212-255, given dc_q(..) and ac_q(..) */
void r1o_quant_ctx_update(r1o_quant_ctx *qc, int tx_size, int is_intra, uint32_t dc_quant, uint32_t ac_quant) {
  qc->log_tx_scale = (uint32_t)r1o_log_tx_scale(tx_size);
  qc->dc_quant = dc_quant;
  r1o_divu_gen(dc_quant, qc->dc_mul_add);
  qc->ac_quant = ac_quant;
  r1o_divu_gen(ac_quant, qc->ac_mul_add);
  qc->dc_offset = (int32_t)dc_quant * (is_intra ? 109 : 108) / 256;
  qc->ac_offset0 = (int32_t)ac_quant * (is_intra ? 98 : 97) / 256;
  qc->ac_offset1 = (int32_t)ac_quant * (is_intra ? 109 : 108) / 256;
  qc->ac_offset_eob = (int32_t)ac_quant * (is_intra ? 88 : 44) / 256;
}

static int32_t signum(int32_t v) { return (v > 0) - (v < 0); }

/* QuantizationContext::quantize, src/quantize.rs:257-316.  `scan` has scan_len entries (the coded
 * area of the transform size: av1_scan_orders[tx_size][tx_type].scan), scan[0] == 0. */
void r1o_quantize(const r1o_quant_ctx *qc, const uint16_t *scan, int scan_len, const int32_t *coeffs, int32_t *qcoeffs) {
  const int lts = (int)qc->log_tx_scale;
  /* (ac_quant as usize - ac_offset_eob as usize).align_power_of_two_and_shift(log_tx_scale) */
  const int32_t deadzone = (int32_t)(((uint64_t)qc->ac_quant - (uint64_t)qc->ac_offset_eob + (1ull << lts) - 1) >> lts);
  /* scan[1..].iter().rposition(|&i| coeffs[i].abs() >= deadzone).map(|n| n + 1).unwrap_or(1) */
  int eob = 1;
  for (int n = scan_len - 2; n >= 0; n--) {
    const int32_t c = coeffs[scan[1 + n]];
    if ((c < 0 ? -c : c) >= deadzone) {
      eob = n + 1;
      break;
    }
  }
  qcoeffs[0] = (int32_t)((uint32_t)coeffs[0] << lts);
  qcoeffs[0] += signum(qcoeffs[0]) * qc->dc_offset;
  qcoeffs[0] = r1o_divu_pair(qcoeffs[0], qc->dc_mul_add);

  int level_mode = 1;
  /* scan[1..].iter().take(eob) */
  for (int k = 0; k < eob && 1 + k < scan_len; k++) {
    const int pos = scan[1 + k];
    const int32_t coeff = (int32_t)((uint32_t)coeffs[pos] << lts);
    const int32_t level0 = r1o_divu_pair(coeff, qc->ac_mul_add);
    const int32_t offset = level0 > 1 - level_mode ? qc->ac_offset1 : qc->ac_offset0;
    const int32_t qcoeff = r1o_divu_pair(coeff + signum(coeff) * offset, qc->ac_mul_add);
    qcoeffs[pos] = qcoeff;
    if (level_mode != 0 && qcoeff == 0) level_mode = 0;
    else if (qcoeff > 1) level_mode = 1;
  }
  /* scan.iter().skip(eob + 1) */
  for (int k = eob + 1; k < scan_len; k++) qcoeffs[scan[k]] = 0;
}

/* dequantize, src/quantize.rs:318-334, over `len` coefficients (the coded area) */
void r1o_dequantize(uint32_t dc_quant, uint32_t ac_quant, int log_tx_scale, const int32_t *qcoeffs, int32_t *rcoeffs,
                    int len) {
  const int32_t offset = (1 << log_tx_scale) - 1;
  for (int i = 0; i < len; i++) {
    const int32_t c = qcoeffs[i];
    const int32_t quant = (int32_t)(i == 0 ? dc_quant : ac_quant);
    rcoeffs[i] = (int32_t)((uint32_t)c * (uint32_t)quant + (uint32_t)((c >> 31) & offset)) >> log_tx_scale;
  }
}

/* n blocks: block i reads coeffs + i * tx_area, writes qcoeffs (and rcoeffs, if given) + i * scan_len */
void r1o_quantize_batch(const r1o_quant_ctx *qc, int tx_size, const uint16_t *scan, const int32_t *coeffs,
                        int32_t *qcoeffs, int32_t *rcoeffs, int64_t n) {
  const int w = r1o_tx_width(tx_size), h = r1o_tx_height(tx_size), area = w * h;
  const int coded = (w < 32 ? w : 32) * (h < 32 ? h : 32);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++) {
    r1o_quantize(qc, scan, coded, coeffs + i * area, qcoeffs + i * coded);
    if (rcoeffs)
      r1o_dequantize(qc->dc_quant, qc->ac_quant, (int)qc->log_tx_scale, qcoeffs + i * coded, rcoeffs + i * coded, coded);
  }
}
