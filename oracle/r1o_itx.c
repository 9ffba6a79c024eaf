/*
 * r1o_itx.c -- CPU ORACLE (test infrastructure only, see r1o.h): inverse transforms.
 *
 * Restates src/transform/inverse.rs:
 *   1-D  av1_idct4 / 8 / 16 / 32 / 64   :35-55, :118-165, :261-356, :541-839, :848-1544
 *        av1_iadst4 / 8 / 16            :63-109, :173-252, :364-532  (+ av1_iflipadst*: reversed output)
 *        av1_iidentity4 / 8 / 16 / 32   :111-116, :254-259, :534-539, :841-846
 *   2-D  NativeInvTxfm2D::inv_txfm2d     :2061-2107, inv_txfm2d_add :2014-2028
 * The reference spells every butterfly out; here the DCT is the recursion those listings unroll
 * (even half = the half-size DCT of the even inputs, odd half = one "primary" rotation stage, then
 * alternating add/sub and rotation levels, then the +/- combine) and the ADST 8 / 16 the loop over
 * halving distances, so that the oracle is not a second copy of the same text.
 * Parity PINNED: tests/test_oracle_vs_reference_src.py runs the reference's own functions (its Rust text
 * transliterated at test time) against these on random inputs, and tests/golden/inv_tx_r01.npz holds
 * vectors made the same way (tests/golden/make_inv_tx_golden.py).
 */
#include <string.h>

#include "r1o.h"

typedef int32_t T;

/* COSPI_INV, inverse.rs:22-29: round(4096 cos(i pi / 128)) */
static const int32_t COSPI[64] = {
    4096, 4095, 4091, 4085, 4076, 4065, 4052, 4036, 4017, 3996, 3973, 3948, 3920, 3889, 3857, 3822,
    3784, 3745, 3703, 3659, 3612, 3564, 3513, 3461, 3406, 3349, 3290, 3229, 3166, 3102, 3035, 2967,
    2896, 2824, 2751, 2675, 2598, 2520, 2440, 2359, 2276, 2191, 2106, 2019, 1931, 1842, 1751, 1660,
    1567, 1474, 1380, 1285, 1189, 1092, 995,  897,  799,  700,  601,  501,  401,  301,  201,  101};
static const int32_t SINPI[5] = {0, 1321, 2482, 3344, 3803}; /* SINPI_INV, :31 */
#define ISQRT2 5793                                           /* SQRT2, src/transform/mod.rs:53 */
#define IINV_SQRT2 2896                                       /* INV_SQRT2, :54 */

/* half_btf, src/transform/mod.rs:478-488: wrapping i32, bit = INV_COS_BIT = 12 */
static inline T btf(int32_t w0, T a, int32_t w1, T b) {
  uint32_t r = (uint32_t)w0 * (uint32_t)a + (uint32_t)w1 * (uint32_t)b + (1u << 11);
  return (T)r >> 12;
}
/* clamp_value, src/transform/mod.rs:492-496 */
static inline T clampv(T v, int bit) {
  const T hi = (T)(((int64_t)1 << (bit - 1)) - 1), lo = (T)(-((int64_t)1 << (bit - 1)));
  return v < lo ? lo : v > hi ? hi : v;
}
static inline T rshift(T v, int bit) { return (v + ((1 << bit) >> 1)) >> bit; } /* src/util/mod.rs:241 */
static inline int brev(int v, int bits) {
  int r = 0;
  for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
  return r;
}
static inline int ilog2(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

/* the odd half of a 2M-point inverse DCT: x[j] = coefficient 2j+1 on entry, the M values the final
 * combine adds / subtracts on exit (idct8 :129-154, idct16 :273-349, idct32 / idct64 stages 2.. on
 * indices >= M) */
static void idct_odd(int M, const T *in, T *x, int range) {
  const int lm = ilog2(M);
  T t[32];
  for (int i = 0; i < M; i++) t[i] = in[brev(i, lm)];
  /* primary rotations: pair (i, M-1-i), angle = odd coefficient index * 64 / 2M */
  for (int i = 0; i < M / 2; i++) {
    const int k = (2 * brev(i, lm) + 1) * 32 / M;
    x[i] = btf(COSPI[64 - k], t[i], -COSPI[k], t[M - 1 - i]);
    x[M - 1 - i] = btf(COSPI[k], t[i], COSPI[64 - k], t[M - 1 - i]);
  }
  for (int g = 2; g <= M / 2; g *= 2) {
    /* add / sub inside groups of g; odd groups mirrored */
    for (int q = 0; q < M / g; q++) {
      T *a = x + q * g;
      for (int i = 0; i < g / 2; i++) {
        const T u = a[i], v = a[g - 1 - i];
        if (q & 1) { a[i] = clampv(-u + v, range); a[g - 1 - i] = clampv(u + v, range); }
        else { a[i] = clampv(u + v, range); a[g - 1 - i] = clampv(u - v, range); }
      }
    }
    /* rotations of the middle half of every block of 2g, pairing i with M-1-i */
    for (int i = 0; i < M / 2; i++) {
      const int m = i % (2 * g);
      if (m < g / 2 || m >= g + g / 2) continue;
      int k = 32;
      if (4 * g <= M) {
        const int nb = M / (4 * g), b = i / (2 * g);
        k = (16 / nb) * (2 * brev(b, ilog2(nb) + 1) + 1);
      }
      const T u = x[i], v = x[M - 1 - i];
      if (m < g) {
        x[i] = btf(-COSPI[k], u, COSPI[64 - k], v);
        x[M - 1 - i] = btf(COSPI[64 - k], u, COSPI[k], v);
      } else {
        x[i] = btf(-COSPI[64 - k], u, -COSPI[k], v);
        x[M - 1 - i] = btf(-COSPI[k], u, COSPI[64 - k], v);
      }
    }
  }
}

static void idct(int n, const T *in, T *out, int range) {
  if (n == 2) { /* the c32 pair of idct4, :44-45 */
    out[0] = btf(COSPI[32], in[0], COSPI[32], in[1]);
    out[1] = btf(COSPI[32], in[0], -COSPI[32], in[1]);
    return;
  }
  const int M = n / 2;
  T ev[32] = {0}, od[32] = {0}, e[32], o[32];
  for (int i = 0; i < M; i++) { ev[i] = in[2 * i]; od[i] = in[2 * i + 1]; }
  idct(M, ev, e, range);
  idct_odd(M, od, o, range); /* M == 2: only the primary rotation, the c48 / c16 pair of :46-47 */
  for (int i = 0; i < M; i++) {
    out[i] = clampv(e[i] + o[M - 1 - i], range);
    out[n - 1 - i] = clampv(e[i] - o[M - 1 - i], range);
  }
}

/* av1_iadst4, :63-109: no clamps, one rounding at the end */
static void iadst4(const T *in, T *out) {
  const T x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3];
  const T a = SINPI[1] * x0 + SINPI[4] * x2 + SINPI[2] * x3; /* s0 after stage 4 */
  const T b = SINPI[2] * x0 - SINPI[1] * x2 - SINPI[4] * x3; /* s1 after stage 4 */
  const T c = SINPI[3] * x1;                                 /* s3 */
  const T d = SINPI[3] * (x0 - x2 + x3);                     /* s2 */
  out[0] = rshift(a + c, 12);
  out[1] = rshift(b + c, 12);
  out[2] = rshift(d, 12);
  out[3] = rshift(a + b - c, 12);
}

/* av1_iadst8 :173-252, av1_iadst16 :364-532 */
static void iadst(int n, const T *in, T *out, int range) {
  const int ln = ilog2(n);
  T x[16];
  for (int j = 0; j < n / 2; j++) { /* stage 1 + 2 */
    const T a = in[n - 1 - 2 * j], b = in[2 * j];
    const int k = (4 * j + 1) * 32 / n;
    x[2 * j] = btf(COSPI[k], a, COSPI[64 - k], b);
    x[2 * j + 1] = btf(COSPI[64 - k], a, -COSPI[k], b);
  }
  for (int h = n / 2; h >= 2; h /= 2) {
    for (int blk = 0; blk < n; blk += 2 * h)
      for (int i = blk; i < blk + h; i++) {
        const T u = x[i], v = x[i + h];
        x[i] = clampv(u + v, range);
        x[i + h] = clampv(u - v, range);
      }
    /* rotate the pairs of the upper half of every block of 2h (of h when h == 2) */
    const int span = h == 2 ? 4 : 2 * h, npair = h == 2 ? 1 : h / 2;
    for (int blk = 0; blk < n; blk += span)
      for (int p = 0; p < npair; p++) {
        const int i = blk + span / 2 + 2 * p, q = npair > 1 ? npair / 2 : 1;
        const int k = h == 2 ? 32 : (4 * (p % q) + 1) * 64 / h;
        const T u = x[i], v = x[i + 1];
        if (p < q) {
          x[i] = btf(COSPI[k], u, COSPI[64 - k], v);
          x[i + 1] = btf(COSPI[64 - k], u, -COSPI[k], v);
        } else {
          x[i] = btf(-COSPI[64 - k], u, COSPI[k], v);
          x[i + 1] = btf(COSPI[k], u, COSPI[64 - k], v);
        }
      }
  }
  for (int i = 0; i < n; i++) { /* last stage: bit-reversed Gray code order, odd outputs negated */
    const T v = x[brev(i ^ (i >> 1), ln)];
    out[i] = (i & 1) ? -v : v;
  }
}

/* type1d: 0 DCT, 1 ADST, 2 FLIPADST, 3 identity; n = 4..64; in / out hold n values (may not alias).
 * returns -1 where the reference has no function (INV_TXFM_FNS, :1610-1634) */
int r1o_inv_txfm1d(int type1d, int n, const int32_t *in, int32_t *out, int range) {
  if (n != 4 && n != 8 && n != 16 && n != 32 && n != 64) return -1;
  switch (type1d) {
    case 0: idct(n, in, out, range); return 0;
    case 1:
    case 2:
      if (n > 16) return -1;
      if (n == 4) iadst4(in, out); else iadst(n, in, out, range);
      if (type1d == 2) /* av1_iflipadst*, :57-60 */
        for (int i = 0; i < n / 2; i++) { const T t = out[i]; out[i] = out[n - 1 - i]; out[n - 1 - i] = t; }
      return 0;
    case 3:
      if (n > 32) return -1;
      for (int i = 0; i < n; i++)
        out[i] = n == 4 ? rshift(ISQRT2 * in[i], 12) : n == 8 ? 2 * in[i]
               : n == 16 ? rshift(ISQRT2 * 2 * in[i], 12) : 4 * in[i];
      return 0;
  }
  return -1;
}

/* ---- 2-D ------------------------------------------------------------------------------------- */
/* column / row 1-D type of a TxType (src/transform/mod.rs:196-213): 0 DCT, 1 ADST, 2 FLIPADST, 3 identity */
static const uint8_t COL1D[R1O_TX_TYPES] = {0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3};
static const uint8_t ROW1D[R1O_TX_TYPES] = {0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2};

static int itx_valid_1d(int t, int n) { return t == 0 ? 1 : t == 3 ? n <= 32 : n <= 16; }

/* the pairs txfm_types::Detail implements (:1583-1607) plus the FLIPADST ones, which the fork's native
 * path leaves `unimplemented!` and which are taken here as av1_iflipadst* (:57-60, :167-170, :358-361) */
int r1o_inv_txfm_valid(int tx_size, int tx_type) {
  if (tx_size < 0 || tx_size >= R1O_TX_SIZES_ALL || tx_type < 0 || tx_type >= R1O_TX_TYPES) return 0;
  return itx_valid_1d(COL1D[tx_type], r1o_tx_height(tx_size)) && itx_valid_1d(ROW1D[tx_type], r1o_tx_width(tx_size));
}

/* InvBlock::INTERMEDIATE_SHIFT, :1656-1665 */
static int intermediate_shift(int tx_size) {
  switch (tx_size) {
    case R1O_TX_4X4: case R1O_TX_4X8: case R1O_TX_8X4: return 0;
    case R1O_TX_16X16: case R1O_TX_16X64: case R1O_TX_64X16: case R1O_TX_64X64:
    case R1O_TX_32X32: case R1O_TX_8X32: case R1O_TX_32X8: return 2;
    default: return 1;
  }
}

/* inv_txfm2d, :2061-2107.  coeffs: min(W,32) * min(H,32) values, row-major with stride min(W,32);
 * residual: W*H values, COLUMN-major (residual[c*H + r]) like the reference's `output.chunks_mut(S::H)` */
int r1o_inv_txfm2d(const int32_t *coeffs, int32_t *residual, int tx_size, int tx_type, int bd) {
  if (!r1o_inv_txfm_valid(tx_size, tx_type) || (bd != 8 && bd != 10 && bd != 12)) return -1;
  const int W = r1o_tx_width(tx_size), H = r1o_tx_height(tx_size);
  const int wi = W < 32 ? W : 32, hi = H < 32 ? H : 32;
  const int rect = ilog2(W) - ilog2(H);
  static _Thread_local T buf[64 * 64];
  T a[64], b[64];
  memset(buf, 0, sizeof(T) * (size_t)W * H);
  int range = bd + 8;
  for (int r = 0; r < hi; r++) {
    memset(a, 0, sizeof(a));
    for (int c = 0; c < wi; c++) {
      T v = coeffs[r * wi + c];
      if (rect == 1 || rect == -1) v = rshift((T)((uint32_t)v * IINV_SQRT2), 12);
      a[c] = clampv(v, range);
    }
    r1o_inv_txfm1d(ROW1D[tx_type], W, a, b, range);
    memcpy(buf + r * W, b, sizeof(T) * (size_t)W);
  }
  range = bd + 6 > 16 ? bd + 6 : 16;
  const int sh = intermediate_shift(tx_size);
  for (int c = 0; c < W; c++) {
    for (int r = 0; r < H; r++) a[r] = clampv(rshift(buf[r * W + c], sh), range);
    r1o_inv_txfm1d(COL1D[tx_type], H, a, b, range);
    for (int r = 0; r < H; r++) residual[c * H + r] = rshift(b[r], 4);
  }
  return 0;
}

/* inv_txfm2d_add, :2014-2028: dst += residual, clamped to [0, 2^bd - 1]; stride in pixels */
int r1o_inv_txfm_add(const int32_t *coeffs, void *dst, ptrdiff_t stride, int px_bytes, int tx_size, int tx_type,
                     int bd) {
  static _Thread_local T res[64 * 64];
  if (r1o_inv_txfm2d(coeffs, res, tx_size, tx_type, bd)) return -1;
  const int W = r1o_tx_width(tx_size), H = r1o_tx_height(tx_size), maxv = (1 << bd) - 1;
  for (int c = 0; c < W; c++)
    for (int r = 0; r < H; r++) {
      int32_t v = px_bytes == 1 ? ((uint8_t *)dst)[r * stride + c] : ((uint16_t *)dst)[r * stride + c];
      v += res[c * H + r];
      v = v < 0 ? 0 : v > maxv ? maxv : v;
      if (px_bytes == 1) ((uint8_t *)dst)[r * stride + c] = (uint8_t)v;
      else ((uint16_t *)dst)[r * stride + c] = (uint16_t)v;
    }
  return 0;
}

/* n blocks of one (tx_size, tx_type): block i reads coeffs + i*coded (coded = min(W,32)*min(H,32)) and is
 * added at (xy[2i], xy[2i+1]) of the plane whose pixel (0,0) is dst0; the blocks must not overlap */
int r1o_inv_txfm_add_batch(const int32_t *coeffs, void *dst0, ptrdiff_t stride, int px_bytes, const int32_t *xy,
                           int64_t n, int tx_size, int tx_type, int bd) {
  if (!r1o_inv_txfm_valid(tx_size, tx_type) || (bd != 8 && bd != 10 && bd != 12)) return -1;
  const int W = r1o_tx_width(tx_size), H = r1o_tx_height(tx_size);
  const int64_t coded = (int64_t)(W < 32 ? W : 32) * (H < 32 ? H : 32);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++)
    r1o_inv_txfm_add(coeffs + i * coded, (uint8_t *)dst0 + ((int64_t)xy[2 * i + 1] * stride + xy[2 * i]) * px_bytes,
                     stride, px_bytes, tx_size, tx_type, bd);
  return 0;
}
