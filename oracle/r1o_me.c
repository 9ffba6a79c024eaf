/*
 * r1o_me.c -- ORACLE (test infrastructure, never shipped): motion-search
 * drivers restated from src/me.rs.
 *   get_mv_range          src/me.rs:64-80
 *   get_mv_rate           src/me.rs:1005-1021
 *   full_search           src/me.rs:943-990 (+ window iterators
 *                         src/tiling/plane_region.rs:175-201,534-600)
 *   get_mv_rd_cost / compute_mv_rd_cost   src/me.rs:787-856
 * Parity: UNPINNED (the reference has no ME tests); literal restatement.
 */
#include "r1o.h"
#include <stdlib.h>

/* src/me.rs:1009-1017 on i16 */
static uint32_t diff_to_rate(int16_t diff, int allow_hp) {
  int16_t d = allow_hp ? diff : (int16_t)(diff >> 1);
  if (d == 0) return 0;
  uint16_t a = (uint16_t)(d < 0 ? -d : d);
  int lz = 0;
  for (int b = 15; b >= 0 && !((a >> b) & 1); b--) lz++;
  return 2u * (16u - (uint32_t)lz);
}

uint32_t r1o_mv_rate(r1o_mv a, r1o_mv b, int allow_hp) {
  return diff_to_rate((int16_t)(a.row - b.row), allow_hp) +
         diff_to_rate((int16_t)(a.col - b.col), allow_hp);
}

/* src/me.rs:64-80 (MI_SIZE = 4) */
void r1o_mv_range(int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h,
                  int64_t out[4]) {
  int64_t border_w = 128 + (int64_t)blk_w * 8;
  int64_t border_h = 128 + (int64_t)blk_h * 8;
  out[0] = -bo_x * (8 * 4) - border_w;
  out[1] = (w_in_b - bo_x - blk_w / 4) * (8 * 4) + border_w;
  out[2] = -bo_y * (8 * 4) - border_h;
  out[3] = (h_in_b - bo_y - blk_h / 4) * (8 * 4) + border_h;
}

static const void *px_ptr(const void *p0, const r1o_plane_cfg *cfg, int64_t x, int64_t y, int pb) {
  return (const uint8_t *)p0 + (y * cfg->stride + x) * pb;
}

/* PlaneRegion::new bounds asserts, src/tiling/plane_region.rs:139-142 */
static int region_ok(const r1o_plane_cfg *c, int64_t x, int64_t y, int64_t w, int64_t h) {
  return x >= -c->xorigin && y >= -c->yorigin && c->xorigin + x + w <= c->stride &&
         c->yorigin + y + h <= c->alloc_height;
}

/* src/me.rs:943-990 */
int r1o_full_search(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                    const r1o_plane_cfg *rcfg, int64_t x_lo, int64_t x_hi, int64_t y_lo,
                    int64_t y_hi, int blk_w, int blk_h, int64_t po_x, int64_t po_y, int step,
                    uint32_t lambda, const r1o_mv pmv[2], int allow_hp, int pb, r1o_mv *best_mv,
                    uint64_t *lowest_cost) {
  if (x_hi < x_lo || y_hi < y_lo) return -1; /* `(x_hi - x_lo) as usize` would wrap */
  if (!region_ok(rcfg, x_lo, y_lo, (x_hi - x_lo) + blk_w, (y_hi - y_lo) + blk_h)) return -1;
  const void *org = px_ptr(org0, ocfg, po_x, po_y, pb);
  /* vert_windows(blk_h).step_by(step) x horz_windows(blk_w).step_by(step):
   * positions y_lo, y_lo+step, ... <= y_hi (raster order, y outer) */
  for (int64_t y = y_lo; y <= y_hi; y += step)
    for (int64_t x = x_lo; x <= x_hi; x += step) {
      uint32_t sad = r1o_sad(org, ocfg->stride, px_ptr(ref0, rcfg, x, y, pb), rcfg->stride, blk_w,
                             blk_h, pb);
      r1o_mv mv;
      mv.row = (int16_t)(8 * (int16_t)((int16_t)y - (int16_t)po_y));
      mv.col = (int16_t)(8 * (int16_t)((int16_t)x - (int16_t)po_x));
      uint32_t r1 = r1o_mv_rate(mv, pmv[0], allow_hp);
      uint32_t r2 = r1o_mv_rate(mv, pmv[1], allow_hp);
      uint32_t rate = r1 < r2 + 1 ? r1 : r2 + 1;
      uint64_t cost = 256ull * sad + (uint64_t)rate * lambda;
      if (cost < *lowest_cost) { /* strict <: first minimum in raster order wins */
        *lowest_cost = cost;
        *best_mv = mv;
      }
    }
  return 0;
}

/* src/me.rs:787-856 with tmp_plane_opt = Some (sub-pel path): predict_inter
 * (single reference, default_filter REGULAR for both axes,
 * src/encoder.rs:705) into a blk-sized plane, then SAD|SATD + rate. */
uint64_t r1o_subpel_cost(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                         const r1o_plane_cfg *rcfg, int64_t po_x, int64_t po_y, int blk_w,
                         int blk_h, r1o_mv cand, const r1o_mv pmv[2], uint32_t lambda,
                         int use_satd, int allow_hp, const int64_t range[4], int bd, int pb) {
  if (cand.col < range[0] || cand.col > range[1]) return UINT64_MAX;
  if (cand.row < range[2] || cand.row > range[3]) return UINT64_MAX;
  int rf, cf;
  int64_t sx, sy;
  r1o_mv_params(rcfg, po_x, po_y, cand.row, cand.col, &rf, &cf, &sx, &sy);
  uint16_t tmp[128 * 128];
  r1o_put_8tap(tmp, blk_w, px_ptr(ref0, rcfg, sx, sy, pb), rcfg->stride, blk_w, blk_h, cf, rf,
               R1O_REGULAR, R1O_REGULAR, bd, pb);
  const void *org = px_ptr(org0, ocfg, po_x, po_y, pb);
  uint32_t d = use_satd ? r1o_satd(org, ocfg->stride, tmp, blk_w, blk_w, blk_h, pb)
                        : r1o_sad(org, ocfg->stride, tmp, blk_w, blk_w, blk_h, pb);
  uint32_t r1 = r1o_mv_rate(cand, pmv[0], allow_hp);
  uint32_t r2 = r1o_mv_rate(cand, pmv[1], allow_hp);
  uint32_t rate = r1 < r2 + 1 ? r1 : r2 + 1;
  return 256ull * d + (uint64_t)rate * lambda;
}

/* get_mv_rd_cost with tmp_plane_opt = None (full-pel path), src/me.rs:794-799,826-836:
 * the reference block sits at po + mv / 8 (division truncating toward zero). */
uint64_t r1o_fullpel_cost(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                          const r1o_plane_cfg *rcfg, int64_t po_x, int64_t po_y, int blk_w,
                          int blk_h, r1o_mv cand, const r1o_mv pmv[2], uint32_t lambda,
                          int use_satd, int allow_hp, const int64_t range[4], int pb) {
  if (cand.col < range[0] || cand.col > range[1]) return UINT64_MAX;
  if (cand.row < range[2] || cand.row > range[3]) return UINT64_MAX;
  const int64_t rx = po_x + cand.col / 8, ry = po_y + cand.row / 8;
  if (!region_ok(rcfg, rx, ry, blk_w, blk_h)) return UINT64_MAX; /* the reference would panic */
  const void *org = px_ptr(org0, ocfg, po_x, po_y, pb);
  const void *rp = px_ptr(ref0, rcfg, rx, ry, pb);
  uint32_t d = use_satd ? r1o_satd(org, ocfg->stride, rp, rcfg->stride, blk_w, blk_h, pb)
                        : r1o_sad(org, ocfg->stride, rp, rcfg->stride, blk_w, blk_h, pb);
  uint32_t r1 = r1o_mv_rate(cand, pmv[0], allow_hp);
  uint32_t r2 = r1o_mv_rate(cand, pmv[1], allow_hp);
  uint32_t rate = r1 < r2 + 1 ? r1 : r2 + 1;
  return 256ull * d + (uint64_t)rate * lambda;
}

/* diamond_me_search, src/me.rs:693-785, with get_best_predictor :655-691.
 * subpixel: radius 4 -> 2 (1 with high-precision MVs), candidates through predict_inter;
 * else radius 16 -> 8 on the full-pel path.  Returns the number of diamond iterations. */
int r1o_diamond_search(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                       const r1o_plane_cfg *rcfg, int64_t po_x, int64_t po_y, int blk_w, int blk_h,
                       const r1o_mv *predictors, int n_pred, const r1o_mv pmv[2], uint32_t lambda,
                       int use_satd, int allow_hp, const int64_t range[4], int subpixel, int bd,
                       int pb, r1o_mv *center_mv, uint64_t *center_cost) {
  static const int16_t pattern[4][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}}; /* (row, col) */
  int16_t radius = subpixel ? 4 : 16;
  const int16_t radius_end = subpixel ? (allow_hp ? 1 : 2) : 8;
#define R1O_COST(mv)                                                                                          \
  (subpixel ? r1o_subpel_cost(org0, ocfg, ref0, rcfg, po_x, po_y, blk_w, blk_h, (mv), pmv, lambda, use_satd, \
                              allow_hp, range, bd, pb)                                                       \
            : r1o_fullpel_cost(org0, ocfg, ref0, rcfg, po_x, po_y, blk_w, blk_h, (mv), pmv, lambda, use_satd, \
                               allow_hp, range, pb))
  r1o_mv c = {0, 0};
  uint64_t cc = UINT64_MAX;
  for (int k = 0; k < n_pred; k++) {
    uint64_t cost = R1O_COST(predictors[k]);
    if (cost < cc) {
      c = predictors[k];
      cc = cost;
    }
  }
  int iters = 0;
  for (;;) {
    uint64_t best = UINT64_MAX;
    r1o_mv best_mv = {0, 0};
    iters++;
    for (int k = 0; k < 4; k++) {
      r1o_mv cand = {(int16_t)(c.row + radius * pattern[k][0]), (int16_t)(c.col + radius * pattern[k][1])};
      uint64_t cost = R1O_COST(cand);
      if (cost < best) {
        best = cost;
        best_mv = cand;
      }
    }
    if (cc <= best) {
      if (radius == radius_end) break;
      radius /= 2;
    } else {
      c = best_mv;
      cc = best;
    }
  }
#undef R1O_COST
  *center_mv = c;
  *center_cost = cc;
  return iters;
}
