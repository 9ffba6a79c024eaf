/*
 * r1o_rdo.c -- CPU ORACLE (test infrastructure only, see r1o.h): RDO distortion.
 *   cdef_dist_wxh_8x8 / cdef_dist_wxh   src/rdo.rs:219-283
 *   sse_wxh                             src/rdo.rs:286-335
 *   compute_mean_importance / _bias     src/rdo.rs:477-508
 *   RawDistortion * f64 -> Distortion   src/rdo.rs:525-530 (truncating, saturating cast)
 *   tx-domain distortion                src/encoder.rs:1210-1224
 * Parity: UNPINNED (the reference has no golden values for these); literal restatement.
 * Built with -ffp-contract=off: the reference's f64 expressions are never fused.
 */
#include <math.h>
#include "r1o.h"

static inline int64_t px_at(const void *p0, int64_t stride, int64_t x, int64_t y, int px_bytes) {
  return px_bytes == 1 ? ((const uint8_t *)p0)[y * stride + x] : ((const uint16_t *)p0)[y * stride + x];
}

/* Rust `f64 as u64`: NaN -> 0, negative -> 0, too large -> u64::MAX */
static inline uint64_t f64_as_u64(double v) {
  if (!(v > 0.0)) return 0;
  if (v >= 18446744073709551616.0) return UINT64_MAX;
  return (uint64_t)v;
}

/* Block dims (MI units) of BlockSize::from_width_and_height(w, h): just w/4, h/4 */

/* src/rdo.rs:477-508.  importances == NULL: bias 1.0 (raw distortion; test aid, not a reference mode). */
double r1o_distortion_bias(const float *importances, int64_t w_in_imp_b, int64_t w_in_b, int64_t h_in_b,
                           int64_t bo_x, int64_t bo_y, int64_t width_mi, int64_t height_mi) {
  if (!importances) return 1.0;
  int64_t x2 = bo_x + width_mi < w_in_b ? bo_x + width_mi : w_in_b;
  int64_t y2 = bo_y + height_mi < h_in_b ? bo_y + height_mi : h_in_b;
  float total = 0.f;
  for (int64_t y = bo_y; y < y2; y++)
    for (int64_t x = bo_x; x < x2; x++) total += importances[y / 2 * w_in_imp_b + x / 2];
  float mean = total / (float)(width_mi * height_mi);
  const float FACTOR = 3.f;
  return (double)(mean / FACTOR) + 0.65;
}

/* sse_wxh, src/rdo.rs:286-335.  (x1,y1)/(x2,y2): region origins in their planes' pixel units;
 * xdec/ydec: src1.plane_cfg decimation (block dims and the frame block offset use it). */
uint64_t r1o_sse_wxh(const void *src1, int64_t stride1, int64_t x1, int64_t y1, const void *src2, int64_t stride2,
                     int64_t x2, int64_t y2, int64_t w, int64_t h, int xdec, int ydec, int px_bytes,
                     const float *importances, int64_t w_in_imp_b, int64_t w_in_b, int64_t h_in_b) {
  const int64_t imp_w = w < 8 ? w : 8, imp_h = h < 8 ? h : 8;
  const int64_t bw = imp_w >> xdec, bh = imp_h >> ydec;
  uint64_t sse = 0;
  for (int64_t by = 0; by < h / bh; by++)
    for (int64_t bx = 0; bx < w / bw; bx++) {
      uint64_t value = 0;
      for (int64_t j = 0; j < bh; j++) {
        uint32_t row = 0;
        for (int64_t i = 0; i < bw; i++) {
          int32_t c = (int16_t)px_at(src1, stride1, x1 + bx * bw + i, y1 + by * bh + j, px_bytes) -
                      (int16_t)px_at(src2, stride2, x2 + bx * bw + i, y2 + by * bh + j, px_bytes);
          row += (uint32_t)(c * c);
        }
        value += row;
      }
      /* frame_block_offset of the sub-region, src/tiling/plane_region.rs:264-276 */
      const int64_t fbx = (x1 + bx * bw) >> (2 - xdec), fby = (y1 + by * bh) >> (2 - ydec);
      const double bias = r1o_distortion_bias(importances, w_in_imp_b, w_in_b, h_in_b, fbx, fby, imp_w >> 2, imp_h >> 2);
      sse += f64_as_u64((double)value * bias);
    }
  return sse;
}

/* cdef_dist_wxh_8x8, src/rdo.rs:219-253 */
static uint64_t cdef_dist_8x8(const void *src1, int64_t stride1, int64_t x1, int64_t y1, const void *src2,
                              int64_t stride2, int64_t x2, int64_t y2, int bit_depth, int px_bytes) {
  const int coeff_shift = bit_depth - 8;
  int32_t sum_s = 0, sum_d = 0;
  int64_t sum_s2 = 0, sum_d2 = 0, sum_sd = 0;
  for (int j = 0; j < 8; j++)
    for (int i = 0; i < 8; i++) {
      int32_t s = (int32_t)px_at(src1, stride1, x1 + i, y1 + j, px_bytes);
      int32_t d = (int32_t)px_at(src2, stride2, x2 + i, y2 + j, px_bytes);
      sum_s += s;
      sum_d += d;
      sum_s2 += (int64_t)(s * s);
      sum_d2 += (int64_t)(d * d);
      sum_sd += (int64_t)(s * d);
    }
  const double svar = (double)(sum_s2 - (((int64_t)sum_s * sum_s + 32) >> 6));
  const double dvar = (double)(sum_d2 - (((int64_t)sum_d * sum_d + 32) >> 6));
  const double sse = (double)(sum_d2 + sum_s2 - 2 * sum_sd);
  const double ssim_boost = (4033.0 / 16384.0) * (svar + dvar + (double)((int64_t)16384 << (2 * coeff_shift))) /
                            sqrt((double)((uint64_t)16265089 << (4 * coeff_shift)) + svar * dvar);
  return f64_as_u64(sse * ssim_boost + 0.5);
}

/* cdef_dist_wxh, src/rdo.rs:256-283 (non-subsampled planes only; bias block = BLOCK_8X8) */
uint64_t r1o_cdef_dist_wxh(const void *src1, int64_t stride1, int64_t x1, int64_t y1, const void *src2,
                           int64_t stride2, int64_t x2, int64_t y2, int64_t w, int64_t h, int bit_depth,
                           int px_bytes, const float *importances, int64_t w_in_imp_b, int64_t w_in_b,
                           int64_t h_in_b) {
  uint64_t sum = 0;
  for (int64_t j = 0; j < h / 8; j++)
    for (int64_t i = 0; i < w / 8; i++) {
      const uint64_t value =
          cdef_dist_8x8(src1, stride1, x1 + i * 8, y1 + j * 8, src2, stride2, x2 + i * 8, y2 + j * 8, bit_depth, px_bytes);
      const double bias =
          r1o_distortion_bias(importances, w_in_imp_b, w_in_b, h_in_b, (x1 + i * 8) >> 2, (y1 + j * 8) >> 2, 2, 2);
      sum += f64_as_u64((double)value * bias);
    }
  return sum;
}

/* tx-domain distortion of one block, src/encoder.rs:1210-1224.  `coeffs` holds tx area values, `rcoeffs`
 * the coded area (64-point dims cut to 32, src/encoder.rs:1143-1155); zip stops at the shorter one. */
uint64_t r1o_tx_dist(const int32_t *coeffs, const int32_t *rcoeffs, int tx_size) {
  const int w = r1o_tx_width(tx_size), h = r1o_tx_height(tx_size);
  const int coded = (w < 32 ? w : 32) * (h < 32 ? h : 32), area = w * h;
  uint64_t t = 0;
  for (int i = 0; i < coded; i++) {
    int32_t c = (int32_t)((uint32_t)coeffs[i] - (uint32_t)rcoeffs[i]);
    t += (uint64_t)(int64_t)(int32_t)((uint32_t)c * (uint32_t)c); /* (c * c) as u64, wrapping i32 product */
  }
  const int log_tx_scale = (area > 256) + (area > 1024); /* get_log_tx_scale, src/quantize.rs:34-39 */
  const int bits = 2 * (3 - log_tx_scale);
  return (t + (1ull << (bits - 1))) >> bits;
}
