/*
 * r1o_dist.c -- ORACLE (test infrastructure, never shipped): SAD / SATD and
 * plane geometry, restated from the reference's scalar ground truth.
 *   get_sad_ref      src/dist.rs:25-46
 *   get_satd_ref     src/dist.rs:197-328
 *   Plane::new       src/frame/plane.rs:215-244
 *   Plane::pad       src/frame/plane.rs:269-314
 *   downsample_from  src/frame/plane.rs:399-423
 * Parity: PINNED (tests/test_oracle_golden.py reproduces the 44 golden values
 * of src/dist.rs:379-402,437-460 and the pad golden of plane.rs:709-752).
 */
#include "r1o.h"
#include <stdlib.h>
#include <string.h>

static const uint8_t BW[R1O_BLOCK_SIZES_ALL] = {4, 4, 8, 8, 8, 16, 16, 16, 32, 32, 32,
                                                64, 64, 64, 128, 128, 4, 16, 8, 32, 16, 64};
static const uint8_t BH[R1O_BLOCK_SIZES_ALL] = {4, 8, 4, 8, 16, 8, 16, 32, 16, 32, 64,
                                                32, 64, 128, 64, 128, 16, 4, 32, 8, 64, 16};

int r1o_block_width(int b) { return (b >= 0 && b < R1O_BLOCK_SIZES_ALL) ? BW[b] : -1; }
int r1o_block_height(int b) { return (b >= 0 && b < R1O_BLOCK_SIZES_ALL) ? BH[b] : -1; }

static inline int32_t px_at(const void *p, ptrdiff_t idx, int px_bytes) {
  return px_bytes == 1 ? (int32_t)((const uint8_t *)p)[idx]
                       : (int32_t)((const uint16_t *)p)[idx];
}

/* src/dist.rs:25-46.  One loop per pixel type (the per-pixel `px_bytes` branch of the first version kept gcc
 * from vectorising it): with -O3 -march=x86-64-v3 the u8 loop compiles to vpsadbw, which makes the timed CPU
 * baseline a fairer stand-in for the reference's AVX2 SAD asm (src/x86/sad_avx.asm) than a scalar loop. */
#include <immintrin.h>
static uint32_t sad_u8(const uint8_t *restrict org, ptrdiff_t os, const uint8_t *restrict ref, ptrdiff_t rs, int w, int h) {
  /* psadbw on 32- / 16- / 8- / 4-byte pieces of a row: the instruction the reference's asm is built on */
  __m256i acc32 = _mm256_setzero_si256();
  __m128i acc = _mm_setzero_si128();
  uint32_t tail = 0;
  for (int r = 0; r < h; r++) {
    const uint8_t *a = org + r * os, *b = ref + r * rs;
    int c = 0;
    for (; c + 32 <= w; c += 32)
      acc32 = _mm256_add_epi64(acc32, _mm256_sad_epu8(_mm256_loadu_si256((const __m256i *)(a + c)), _mm256_loadu_si256((const __m256i *)(b + c))));
    for (; c + 16 <= w; c += 16)
      acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_loadu_si128((const __m128i *)(a + c)), _mm_loadu_si128((const __m128i *)(b + c))));
    for (; c + 8 <= w; c += 8)
      acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_loadl_epi64((const __m128i *)(a + c)), _mm_loadl_epi64((const __m128i *)(b + c))));
    for (; c + 4 <= w; c += 4) {
      int32_t x, y;
      memcpy(&x, a + c, 4);
      memcpy(&y, b + c, 4);
      acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_cvtsi32_si128(x), _mm_cvtsi32_si128(y)));
    }
    for (; c < w; c++) tail += (uint32_t)(a[c] > b[c] ? a[c] - b[c] : b[c] - a[c]);
  }
  acc = _mm_add_epi64(acc, _mm_add_epi64(_mm256_castsi256_si128(acc32), _mm256_extracti128_si256(acc32, 1)));
  return tail + (uint32_t)_mm_cvtsi128_si32(acc) + (uint32_t)_mm_extract_epi32(acc, 2);
}
static uint32_t sad_u16(const uint16_t *restrict org, ptrdiff_t os, const uint16_t *restrict ref, ptrdiff_t rs, int w, int h) {
  uint32_t sum = 0;
  for (int r = 0; r < h; r++) {
    const uint16_t *a = org + r * os, *b = ref + r * rs;
    uint32_t row = 0;
    for (int c = 0; c < w; c++) row += (uint32_t)(a[c] > b[c] ? a[c] - b[c] : b[c] - a[c]);
    sum += row;
  }
  return sum;
}
uint32_t r1o_sad(const void *org, ptrdiff_t os, const void *ref, ptrdiff_t rs,
                 int w, int h, int px_bytes) {
  return px_bytes == 1 ? sad_u8((const uint8_t *)org, os, (const uint8_t *)ref, rs, w, h)
                       : sad_u16((const uint16_t *)org, os, (const uint16_t *)ref, rs, w, h);
}

/* src/dist.rs:208-256: 1-D Hadamard over n lines; stride0 between lines,
 * stride1 between the elements of a line. */
static void hadamard4_1d(int32_t *data, int n, int stride0, int stride1) {
  for (int i = 0; i < n; i++) {
    int32_t *s = data + i * stride0;
    int32_t a0 = s[0] + s[stride1], a1 = s[0] - s[stride1];
    int32_t a2 = s[2 * stride1] + s[3 * stride1], a3 = s[2 * stride1] - s[3 * stride1];
    s[0] = a0 + a2;
    s[stride1] = a1 + a3;
    s[2 * stride1] = a0 - a2;
    s[3 * stride1] = a1 - a3;
  }
}

static void hadamard8_1d(int32_t *data, int n, int stride0, int stride1) {
  for (int i = 0; i < n; i++) {
    int32_t *s = data + i * stride0;
    int32_t a[8], b[8];
    for (int k = 0; k < 4; k++) {
      a[2 * k] = s[(2 * k) * stride1] + s[(2 * k + 1) * stride1];
      a[2 * k + 1] = s[(2 * k) * stride1] - s[(2 * k + 1) * stride1];
    }
    /* (b0,b2)=bf(a0,a2) (b1,b3)=bf(a1,a3) (b4,b6)=bf(a4,a6) (b5,b7)=bf(a5,a7) */
    b[0] = a[0] + a[2]; b[2] = a[0] - a[2];
    b[1] = a[1] + a[3]; b[3] = a[1] - a[3];
    b[4] = a[4] + a[6]; b[6] = a[4] - a[6];
    b[5] = a[5] + a[7]; b[7] = a[5] - a[7];
    /* (c0,c4)=bf(b0,b4) ... (c3,c7)=bf(b3,b7) */
    for (int k = 0; k < 4; k++) {
      s[k * stride1] = b[k] + b[k + 4];
      s[(k + 4) * stride1] = b[k] - b[k + 4];
    }
  }
}

/* src/dist.rs:282-328 */
uint32_t r1o_satd(const void *org, ptrdiff_t os, const void *ref, ptrdiff_t rs,
                  int w, int h, int px_bytes) {
  int size = w < h ? w : h;
  if (size > 8) size = 8;
  uint64_t sum = 0;
  for (int cy = 0; cy < h; cy += size)
    for (int cx = 0; cx < w; cx += size) {
      int32_t buf[64];
      for (int r = 0; r < size; r++)
        for (int c = 0; c < size; c++)
          buf[r * size + c] = px_at(org, (cy + r) * os + cx + c, px_bytes) -
                              px_at(ref, (cy + r) * rs + cx + c, px_bytes);
      /* hadamard2d: vertical (n=w lines, stride0=1, stride1=h) then
       * horizontal (n=h lines, stride0=w, stride1=1); w==h==size here. */
      if (size == 4) {
        hadamard4_1d(buf, 4, 1, 4);
        hadamard4_1d(buf, 4, 4, 1);
      } else {
        hadamard8_1d(buf, 8, 1, 8);
        hadamard8_1d(buf, 8, 8, 1);
      }
      for (int i = 0; i < size * size; i++)
        sum += (uint64_t)(buf[i] < 0 ? -(int64_t)buf[i] : (int64_t)buf[i]);
    }
  int ln = size == 4 ? 2 : 3; /* msb(size) */
  return (uint32_t)((sum + ((1ull << ln) >> 1)) >> ln);
}

/* The fork's GENERATED SATD kernels compute in i16 lanes for every pixel type (build/kernel/gen/dist.rs:189-194 and
 * :252-257: `let calc_prim = PrimType::I16`, the per-type choice is left unused), so 10- and 12-bit input of high
 * amplitude wraps where get_satd_ref (above) does not.  This restates that arithmetic (satd_4nx4n_kernel :171-241,
 * satd_8nx8n_kernel :242-350, normalisation :553-556) for the `compat_i16_wrap` switch of SURVEY 7 / A16:
 *   - every add / sub of the butterflies wraps in 16 bits; +,- mod 2^16 form a ring, so a lane holds the exact
 *     Hadamard term reduced mod 2^16 whatever the order of the stages;
 *   - the last butterfly of the second dimension is merged with the abs: (|a+b| + |a-b|) = 2 max(|a|, |b|), computed as
 *     `a.abs().max(b.abs()) * 2` in i16 (abs and the doubling wrap), on the pairs the generated code forms: lanes k and
 *     k + 4 of the two half-transforms (8x8: 4-point transforms of columns 0-3 and 4-7; 4x4: c0 +- c1 against c2 +- c3);
 *   - each i16 lane is cast to u32 (`from_cast`: sign extension) and summed with wrapping u32 adds.
 * Unpinned: the generator emits code for a SIMD crate that is not part of /root/reference, so this text cannot be
 * executed here; it is test infrastructure for a later comparison with a real reference binary. */
static int16_t w16(int32_t v) { return (int16_t)(uint16_t)(uint32_t)v; }
static int16_t abs16(int16_t v) { return w16(v < 0 ? -(int32_t)v : (int32_t)v); }
static uint32_t merged16(int32_t a, int32_t b) {
  int16_t x = abs16(w16(a)), y = abs16(w16(b));
  int16_t m = x > y ? x : y;
  return (uint32_t)(int32_t)w16(2 * (int32_t)m);
}

uint32_t r1o_satd_i16(const void *org, ptrdiff_t os, const void *ref, ptrdiff_t rs, int w, int h, int px_bytes) {
  int size = w < h ? w : h;
  if (size > 8) size = 8;
  uint32_t sum = 0;
  for (int cy = 0; cy < h; cy += size)
    for (int cx = 0; cx < w; cx += size) {
      int32_t buf[64];
      for (int r = 0; r < size; r++)
        for (int c = 0; c < size; c++)
          buf[r * size + c] = px_at(org, (cy + r) * os + cx + c, px_bytes) - px_at(ref, (cy + r) * rs + cx + c, px_bytes);
      if (size == 4) {
        hadamard4_1d(buf, 4, 1, 4); /* vertical, exact */
        for (int k = 0; k < 4; k++) {
          const int32_t *c = buf + 4 * k;
          sum += merged16(c[0] + c[1], c[2] + c[3]);
          sum += merged16(c[0] - c[1], c[2] - c[3]);
        }
      } else {
        hadamard8_1d(buf, 8, 1, 8);
        for (int j = 0; j < 8; j++) {
          int32_t *c = buf + 8 * j;
          hadamard4_1d(c, 1, 4, 1);     /* columns 0-3 of this vertical frequency */
          hadamard4_1d(c + 4, 1, 4, 1); /* columns 4-7 */
          for (int k = 0; k < 4; k++) sum += merged16(c[k], c[4 + k]);
        }
      }
    }
  int ln = size == 4 ? 2 : 3;
  return (sum + ((1u << ln) >> 1)) >> ln;
}

/* ---- plane geometry ----------------------------------------------------- */
static int64_t align_pow2(int64_t v, int n) { return (v + (1ll << n) - 1) & ~((1ll << n) - 1); }

/* src/frame/plane.rs:215-244 (STRIDE_ALIGNMENT_LOG2 = 5) */
void r1o_plane_cfg_new(r1o_plane_cfg *cfg, int64_t width, int64_t height, int64_t xdec,
                       int64_t ydec, int64_t xpad, int64_t ypad, int px_bytes) {
  int al = 5 + 1 - px_bytes;
  cfg->xorigin = align_pow2(xpad, al);
  cfg->yorigin = ypad;
  cfg->stride = align_pow2(cfg->xorigin + width + xpad, al);
  cfg->alloc_height = cfg->yorigin + height + ypad;
  cfg->width = width;
  cfg->height = height;
  cfg->xdec = xdec;
  cfg->ydec = ydec;
  cfg->xpad = xpad;
  cfg->ypad = ypad;
}

/* src/frame/plane.rs:269-314 */
void r1o_plane_pad(void *data, const r1o_plane_cfg *cfg, int64_t w, int64_t h, int px_bytes) {
  int64_t xo = cfg->xorigin, yo = cfg->yorigin, stride = cfg->stride;
  int64_t width = (w + cfg->xdec) >> cfg->xdec;
  int64_t height = (h + cfg->ydec) >> cfg->ydec;
  uint8_t *d8 = (uint8_t *)data;
  uint16_t *d16 = (uint16_t *)data;
  if (xo > 0)
    for (int64_t y = 0; y < height; y++) {
      int64_t base = (yo + y) * stride;
      for (int64_t x = 0; x < xo; x++) {
        if (px_bytes == 1) d8[base + x] = d8[base + xo];
        else d16[base + x] = d16[base + xo];
      }
    }
  if (xo + width < stride)
    for (int64_t y = 0; y < height; y++) {
      int64_t base = (yo + y) * stride + xo + width;
      for (int64_t x = 0; x < stride - (xo + width); x++) {
        if (px_bytes == 1) d8[base + x] = d8[base - 1];
        else d16[base + x] = d16[base - 1];
      }
    }
  size_t rowb = (size_t)stride * px_bytes;
  if (yo > 0)
    for (int64_t y = 0; y < yo; y++) memcpy(d8 + y * rowb, d8 + yo * rowb, rowb);
  if (yo + height < cfg->alloc_height)
    for (int64_t y = yo + height; y < cfg->alloc_height; y++)
      memcpy(d8 + y * rowb, d8 + (yo + height - 1) * rowb, rowb);
}

/* src/frame/plane.rs:399-423 */
int r1o_plane_downsample(void *dst, const r1o_plane_cfg *dc, const void *src,
                         const r1o_plane_cfg *sc, int px_bytes) {
  if (dc->width * 2 != sc->width || dc->height * 2 != sc->height) return -1;
  for (int64_t row = 0; row < dc->height; row++)
    for (int64_t col = 0; col < dc->width; col++) {
      int64_t s0 = (sc->yorigin + 2 * row) * sc->stride + sc->xorigin + 2 * col;
      uint32_t sum = (uint32_t)px_at(src, s0, px_bytes) + (uint32_t)px_at(src, s0 + 1, px_bytes) +
                     (uint32_t)px_at(src, s0 + sc->stride, px_bytes) +
                     (uint32_t)px_at(src, s0 + sc->stride + 1, px_bytes);
      uint32_t avg = (sum + 2) >> 2;
      int64_t di = (dc->yorigin + row) * dc->stride + dc->xorigin + col;
      if (px_bytes == 1) ((uint8_t *)dst)[di] = (uint8_t)avg;
      else ((uint16_t *)dst)[di] = (uint16_t)avg;
    }
  return 0;
}
