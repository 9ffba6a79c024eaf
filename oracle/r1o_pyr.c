/*
 * r1o_pyr.c -- CPU ORACLE (test infrastructure only, see r1o.h): the full-pel stages of FullSearch
 * motion estimation on the frame pyramid, one block per call.
 *   adjust_bo                 src/me.rs:993-1003
 *   estimate_motion_ss4       src/me.rs:1023-1075
 *   estimate_motion_ss2       src/me.rs:280-333  + FullSearch::me_ss2        :597-650
 *   estimate_motion           src/me.rs:335-385  + FullSearch::full_pixel_me :522-565
 * Parity: UNPINNED (no reference vectors); literal restatement on top of r1o_full_search.
 */
#include "r1o.h"

static void adjust_bo(int64_t bo_x, int64_t bo_y, int64_t mi_w, int64_t mi_h, int blk_w, int blk_h, int64_t *ax, int64_t *ay) {
  int64_t x = bo_x < mi_w - blk_w / 4 ? bo_x : mi_w - blk_w / 4;
  int64_t y = bo_y < mi_h - blk_h / 4 ? bo_y : mi_h - blk_h / 4;
  *ax = x > 0 ? x : 0;
  *ay = y > 0 ? y : 0;
}
static int64_t mx(int64_t a, int64_t b) { return a > b ? a : b; }
static int64_t mn(int64_t a, int64_t b) { return a < b ? a : b; }

int r1o_estimate_motion_ss4(const void *org_q, const r1o_plane_cfg *ocfg, const void *ref_q, const r1o_plane_cfg *rcfg,
                            int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h,
                            int me_range_scale, double me_lambda, int allow_hp, int pb, r1o_mv *out) {
  int64_t ax, ay, r[4];
  adjust_bo(bo_x, bo_y, w_in_b, h_in_b, blk_w, blk_h, &ax, &ay);
  const int64_t po_x = (ax << 2) >> 2, po_y = (ay << 2) >> 2;
  const int64_t range_x = 192 * (int64_t)me_range_scale, range_y = 64 * (int64_t)me_range_scale;
  r1o_mv_range(w_in_b, h_in_b, ax, ay, blk_w, blk_h, r);
  const int64_t x_lo = po_x + (mx(-range_x, r[0] / 8) >> 2), x_hi = po_x + (mn(range_x, r[1] / 8) >> 2);
  const int64_t y_lo = po_y + (mx(-range_y, r[2] / 8) >> 2), y_hi = po_y + (mn(range_y, r[3] / 8) >> 2);
  const uint32_t lambda = (uint32_t)(me_lambda * 256.0 / 16.0 * 0.125);
  r1o_mv best = {0, 0}, zero[2] = {{0, 0}, {0, 0}};
  uint64_t cost = UINT64_MAX;
  int rc = r1o_full_search(org_q, ocfg, ref_q, rcfg, x_lo, x_hi, y_lo, y_hi, blk_w >> 2, blk_h >> 2, po_x, po_y, 1, lambda, zero,
                           allow_hp, pb, &best, &cost);
  out->row = (int16_t)(best.row * 4);
  out->col = (int16_t)(best.col * 4);
  return rc;
}

/* shift = 1, step = 1: me_ss2 on half-res planes; shift = 0, step = 2: full_pixel_me */
static int rounds(const void *org, const r1o_plane_cfg *ocfg, const void *ref, const r1o_plane_cfg *rcfg, int64_t w_in_b,
                  int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h, const r1o_mv *preds, const uint8_t *valid,
                  int slots, int shift, int step, uint32_t lambda, int allow_hp, int pb, r1o_mv *best) {
  int64_t ax, ay, r[4];
  adjust_bo(bo_x, bo_y, w_in_b, h_in_b, blk_w, blk_h, &ax, &ay);
  const int64_t po_x = (ax << 2) >> shift, po_y = (ay << 2) >> shift;
  r1o_mv_range(w_in_b, h_in_b, ax, ay, blk_w, blk_h, r);
  r1o_mv zero[2] = {{0, 0}, {0, 0}};
  uint64_t cost = UINT64_MAX;
  best->row = best->col = 0;
  for (int s = 0; s < slots; s++) {
    if (!valid[s]) continue;
    const int64_t range = 16, pc = preds[s].col / 8, pr = preds[s].row / 8;
    const int64_t x_lo = po_x + (mn(mx(pc - range, r[0] / 8), r[1] / 8) >> shift);
    const int64_t x_hi = po_x + (mn(mx(pc + range, r[0] / 8), r[1] / 8) >> shift);
    const int64_t y_lo = po_y + (mn(mx(pr - range, r[2] / 8), r[3] / 8) >> shift);
    const int64_t y_hi = po_y + (mn(mx(pr + range, r[2] / 8), r[3] / 8) >> shift);
    int rc = r1o_full_search(org, ocfg, ref, rcfg, x_lo, x_hi, y_lo, y_hi, blk_w >> shift, blk_h >> shift, po_x, po_y, step, lambda,
                             zero, allow_hp, pb, best, &cost);
    if (rc) return rc;
  }
  return 0;
}

int r1o_estimate_motion_ss2(const void *org_h, const r1o_plane_cfg *ocfg, const void *ref_h, const r1o_plane_cfg *rcfg,
                            int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h,
                            const r1o_mv pmvs[3], const uint8_t valid[3], double me_lambda, int allow_hp, int pb, r1o_mv *out) {
  r1o_mv best;
  int rc = rounds(org_h, ocfg, ref_h, rcfg, w_in_b, h_in_b, bo_x, bo_y, blk_w, blk_h, pmvs, valid, 3, 1, 1,
                  (uint32_t)(me_lambda * 256.0 / 4.0 * 0.125), allow_hp, pb, &best);
  out->row = (int16_t)(best.row * 2);
  out->col = (int16_t)(best.col * 2);
  return rc;
}

int r1o_estimate_motion(const void *org, const r1o_plane_cfg *ocfg, const void *ref, const r1o_plane_cfg *rcfg, int64_t w_in_b,
                        int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w, int blk_h, const r1o_mv cmvs[7],
                        const uint8_t valid[7], double me_lambda, int allow_hp, int pb, r1o_mv *out) {
  return rounds(org, ocfg, ref, rcfg, w_in_b, h_in_b, bo_x, bo_y, blk_w, blk_h, cmvs, valid, 7, 0, 2,
                (uint32_t)(me_lambda * 256.0 * 0.5), allow_hp, pb, out);
}
