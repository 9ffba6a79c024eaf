/*
 * r1o_mc.c -- ORACLE (test infrastructure, never shipped): sub-pel motion
 * compensation restated from the reference's scalar `_ref` functions.
 *   SUBPEL_FILTERS   src/mc.rs:70-179
 *   get_filter       src/mc.rs:201-210
 *   put_8tap_ref     src/mc.rs:213-307
 *   prep_8tap_ref    src/mc.rs:310-387
 *   mc_avg_ref       src/mc.rs:389-408
 *   get_params       src/predict.rs:267-284 (+ PlaneSlice::clamp plane.rs:521-533)
 * Parity: UNPINNED -- the fork has no MC test vectors; this
 * is a literal restatement, cross-checked in tests/test_oracle_props.py
 * against the AV1-spec rounding (InterRound0/1) and copy/shift identities.
 */
#include "r1o.h"

/* [bank][frac][tap]; banks 0..3 = REGULAR, SMOOTH, SHARP, BILINEAR;
 * 4,5 = the 4-tap REGULAR / SMOOTH variants used when length <= 4. */
static const int16_t FILT[6][16][8] = {
    {{0, 0, 0, 128, 0, 0, 0, 0},
     {0, 2, -6, 126, 8, -2, 0, 0},     {0, 2, -10, 122, 18, -4, 0, 0},
     {0, 2, -12, 116, 28, -8, 2, 0},   {0, 2, -14, 110, 38, -10, 2, 0},
     {0, 2, -14, 102, 48, -12, 2, 0},  {0, 2, -16, 94, 58, -12, 2, 0},
     {0, 2, -14, 84, 66, -12, 2, 0},   {0, 2, -14, 76, 76, -14, 2, 0},
     {0, 2, -12, 66, 84, -14, 2, 0},   {0, 2, -12, 58, 94, -16, 2, 0},
     {0, 2, -12, 48, 102, -14, 2, 0},  {0, 2, -10, 38, 110, -14, 2, 0},
     {0, 2, -8, 28, 116, -12, 2, 0},   {0, 0, -4, 18, 122, -10, 2, 0},
     {0, 0, -2, 8, 126, -6, 2, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0},      {0, 2, 28, 62, 34, 2, 0, 0},
     {0, 0, 26, 62, 36, 4, 0, 0},      {0, 0, 22, 62, 40, 4, 0, 0},
     {0, 0, 20, 60, 42, 6, 0, 0},      {0, 0, 18, 58, 44, 8, 0, 0},
     {0, 0, 16, 56, 46, 10, 0, 0},     {0, -2, 16, 54, 48, 12, 0, 0},
     {0, -2, 14, 52, 52, 14, -2, 0},   {0, 0, 12, 48, 54, 16, -2, 0},
     {0, 0, 10, 46, 56, 16, 0, 0},     {0, 0, 8, 44, 58, 18, 0, 0},
     {0, 0, 6, 42, 60, 20, 0, 0},      {0, 0, 4, 40, 62, 22, 0, 0},
     {0, 0, 4, 36, 62, 26, 0, 0},      {0, 0, 2, 34, 62, 28, 2, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0},      {-2, 2, -6, 126, 8, -2, 2, 0},
     {-2, 6, -12, 124, 16, -6, 4, -2}, {-2, 8, -18, 120, 26, -10, 6, -2},
     {-4, 10, -22, 116, 38, -14, 6, -2}, {-4, 10, -22, 108, 48, -18, 8, -2},
     {-4, 10, -24, 100, 60, -20, 8, -2}, {-4, 10, -24, 90, 70, -22, 10, -2},
     {-4, 12, -24, 80, 80, -24, 12, -4}, {-2, 10, -22, 70, 90, -24, 10, -4},
     {-2, 8, -20, 60, 100, -24, 10, -4}, {-2, 8, -18, 48, 108, -22, 10, -4},
     {-2, 6, -14, 38, 116, -22, 10, -4}, {-2, 6, -10, 26, 120, -18, 8, -2},
     {-2, 4, -6, 16, 124, -12, 6, -2}, {0, 2, -2, 8, 126, -6, 2, -2}},
    {{0, 0, 0, 128, 0, 0, 0, 0},      {0, 0, 0, 120, 8, 0, 0, 0},
     {0, 0, 0, 112, 16, 0, 0, 0},      {0, 0, 0, 104, 24, 0, 0, 0},
     {0, 0, 0, 96, 32, 0, 0, 0},       {0, 0, 0, 88, 40, 0, 0, 0},
     {0, 0, 0, 80, 48, 0, 0, 0},       {0, 0, 0, 72, 56, 0, 0, 0},
     {0, 0, 0, 64, 64, 0, 0, 0},       {0, 0, 0, 56, 72, 0, 0, 0},
     {0, 0, 0, 48, 80, 0, 0, 0},       {0, 0, 0, 40, 88, 0, 0, 0},
     {0, 0, 0, 32, 96, 0, 0, 0},       {0, 0, 0, 24, 104, 0, 0, 0},
     {0, 0, 0, 16, 112, 0, 0, 0},      {0, 0, 0, 8, 120, 0, 0, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0},      {0, 0, -4, 126, 8, -2, 0, 0},
     {0, 0, -8, 122, 18, -4, 0, 0},    {0, 0, -10, 116, 28, -6, 0, 0},
     {0, 0, -12, 110, 38, -8, 0, 0},   {0, 0, -12, 102, 48, -10, 0, 0},
     {0, 0, -14, 94, 58, -10, 0, 0},   {0, 0, -12, 84, 66, -10, 0, 0},
     {0, 0, -12, 76, 76, -12, 0, 0},   {0, 0, -10, 66, 84, -12, 0, 0},
     {0, 0, -10, 58, 94, -14, 0, 0},   {0, 0, -10, 48, 102, -12, 0, 0},
     {0, 0, -8, 38, 110, -12, 0, 0},   {0, 0, -6, 28, 116, -10, 0, 0},
     {0, 0, -4, 18, 122, -8, 0, 0},    {0, 0, -2, 8, 126, -4, 0, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0},      {0, 0, 30, 62, 34, 2, 0, 0},
     {0, 0, 26, 62, 36, 4, 0, 0},      {0, 0, 22, 62, 40, 4, 0, 0},
     {0, 0, 20, 60, 42, 6, 0, 0},      {0, 0, 18, 58, 44, 8, 0, 0},
     {0, 0, 16, 56, 46, 10, 0, 0},     {0, 0, 14, 54, 48, 12, 0, 0},
     {0, 0, 12, 52, 52, 12, 0, 0},     {0, 0, 12, 48, 54, 14, 0, 0},
     {0, 0, 10, 46, 56, 16, 0, 0},     {0, 0, 8, 44, 58, 18, 0, 0},
     {0, 0, 6, 42, 60, 20, 0, 0},      {0, 0, 4, 40, 62, 22, 0, 0},
     {0, 0, 4, 36, 62, 26, 0, 0},      {0, 0, 2, 34, 62, 30, 0, 0}},
};

static inline int32_t tap(int bank, int frac, int k) { return FILT[bank][frac][k]; }

/* src/mc.rs:201-210 */
static void get_filter(int32_t f[8], int mode, int frac, int length) {
  int idx = (mode == R1O_BILINEAR || length > 4) ? mode : ((mode < 1 ? mode : 1) + 4);
  for (int k = 0; k < 8; k++) f[k] = tap(idx, frac, k);
}

static inline int32_t rshift(int32_t v, int bit) { return (v + ((1 << bit) >> 1)) >> bit; }
static inline int32_t clampi(int32_t v, int32_t lo, int32_t hi) { return v < lo ? lo : (v > hi ? hi : v); }

static inline int32_t ld(const void *p, ptrdiff_t i, int pb) {
  return pb == 1 ? (int32_t)((const uint8_t *)p)[i] : (int32_t)((const uint16_t *)p)[i];
}
static inline void st(void *p, ptrdiff_t i, int32_t v, int pb) {
  if (pb == 1) ((uint8_t *)p)[i] = (uint8_t)v;
  else ((uint16_t *)p)[i] = (uint16_t)v;
}
/* run_filter src/mc.rs:187-198 over pixels */
static inline int32_t filt_px(const void *p, ptrdiff_t i, ptrdiff_t step, const int32_t f[8], int pb) {
  int32_t s = 0;
  for (int k = 0; k < 8; k++) s += f[k] * ld(p, i + k * step, pb);
  return s;
}
static inline int32_t filt_i16(const int16_t *p, ptrdiff_t step, const int32_t f[8]) {
  int32_t s = 0;
  for (int k = 0; k < 8; k++) s += f[k] * (int32_t)p[k * step];
  return s;
}

/* Shared 4-case engine.  mode_out: 0 = put (clamped pixels), 1 = prep (i16). */
static void mc_engine(int is_prep, void *dst, ptrdiff_t ds, int16_t *tmp, const void *src,
                      ptrdiff_t ss, int w, int h, int cf, int rf, int mx, int my, int bd, int pb) {
  int32_t yf[8], xf[8];
  get_filter(yf, my, rf, h);
  get_filter(xf, mx, cf, w);
  const int32_t maxv = (1 << bd) - 1;
  const int ib = 4 - (bd == 12 ? 2 : 0);
  if (cf == 0 && rf == 0) {
    for (int r = 0; r < h; r++)
      for (int c = 0; c < w; c++) {
        int32_t p = ld(src, r * ss + c, pb);
        if (is_prep) tmp[r * w + c] = (int16_t)(p << ib);     /* mc.rs:324-325 */
        else st(dst, r * ds + c, p, pb);                       /* mc.rs:225-229 */
      }
  } else if (cf == 0) { /* vertical only: rows -3..+4 */
    for (int r = 0; r < h; r++)
      for (int c = 0; c < w; c++) {
        int32_t s = filt_px(src, (r - 3) * ss + c, ss, yf, pb);
        if (is_prep) tmp[r * w + c] = (int16_t)rshift(s, 7 - ib);          /* mc.rs:334-339 */
        else st(dst, r * ds + c, clampi(rshift(s, 7), 0, maxv), pb);        /* mc.rs:237-246 */
      }
  } else if (rf == 0) { /* horizontal only */
    for (int r = 0; r < h; r++)
      for (int c = 0; c < w; c++) {
        int32_t s = filt_px(src, r * ss + c - 3, 1, xf, pb);
        if (is_prep) tmp[r * w + c] = (int16_t)rshift(s, 7 - ib);          /* mc.rs:348-351 */
        else st(dst, r * ds + c, clampi(rshift(rshift(s, 7 - ib), ib), 0, maxv), pb); /* :256-266 */
      }
  } else { /* 2-D: H pass into i16 (truncating cast), then V pass */
    int16_t inter[8 * (128 + 7)];
    for (int cg = 0; cg < w; cg += 8) {
      int cend = cg + 8 < w ? cg + 8 : w;
      for (int r = 0; r < h + 7; r++)
        for (int c = cg; c < cend; c++)
          inter[8 * r + (c - cg)] =
              (int16_t)rshift(filt_px(src, (r - 3) * ss + c - 3, 1, xf, pb), 7 - ib); /* :278-281 */
      for (int r = 0; r < h; r++)
        for (int c = cg; c < cend; c++) {
          int32_t s = filt_i16(inter + 8 * r + (c - cg), 8, yf);
          if (is_prep) tmp[r * w + c] = (int16_t)rshift(s, 7);                       /* :372-381 */
          else st(dst, r * ds + c, clampi(rshift(s, 7 + ib), 0, maxv), pb);           /* :288-301 */
        }
    }
  }
}

void r1o_put_8tap(void *dst, ptrdiff_t ds, const void *src, ptrdiff_t ss, int w, int h, int cf,
                  int rf, int mx, int my, int bd, int pb) {
  mc_engine(0, dst, ds, 0, src, ss, w, h, cf, rf, mx, my, bd, pb);
}

void r1o_prep_8tap(int16_t *tmp, const void *src, ptrdiff_t ss, int w, int h, int cf, int rf,
                   int mx, int my, int bd, int pb) {
  mc_engine(1, 0, 0, tmp, src, ss, w, h, cf, rf, mx, my, bd, pb);
}

/* src/mc.rs:389-408 */
void r1o_mc_avg(void *dst, ptrdiff_t ds, const int16_t *t1, const int16_t *t2, int w, int h,
                int bd, int pb) {
  const int32_t maxv = (1 << bd) - 1;
  const int ib = 4 - (bd == 12 ? 2 : 0);
  for (int r = 0; r < h; r++)
    for (int c = 0; c < w; c++)
      st(dst, r * ds + c,
         clampi(rshift((int32_t)t1[r * w + c] + (int32_t)t2[r * w + c], ib + 1), 0, maxv), pb);
}

/* src/predict.rs:267-284 */
void r1o_mv_params(const r1o_plane_cfg *cfg, int64_t po_x, int64_t po_y, int mv_row, int mv_col,
                   int *row_frac, int *col_frac, int64_t *src_x, int64_t *src_y) {
  int sr = 3 + (int)cfg->ydec, sc = 3 + (int)cfg->xdec;
  int32_t ro = (int32_t)mv_row >> sr, co = (int32_t)mv_col >> sc;
  *row_frac = (mv_row - (ro << sr)) << (4 - sr);
  *col_frac = (mv_col - (co << sc)) << (4 - sc);
  int64_t qx = po_x + co - 3, qy = po_y + ro - 3;
  /* PlaneSlice::clamp: min(width) then max(-origin) */
  if (qx > cfg->width) qx = cfg->width;
  if (qx < -cfg->xorigin) qx = -cfg->xorigin;
  if (qy > cfg->height) qy = cfg->height;
  if (qy < -cfg->yorigin) qy = -cfg->yorigin;
  *src_x = qx + 3;
  *src_y = qy + 3;
}
