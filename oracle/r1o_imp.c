/*
 * r1o_imp.c -- CPU ORACLE (test infrastructure only, see r1o.h): block importances.
 *
 * Restates the two loops of ContextInner::compute_block_importances
 * (src/api/internal.rs:823-1075):
 *   r1o_importance_propagate   :897-1047  one (frame, reference) pair: inter cost = get_satd of every
 *                              8x8 importance block against the reference at its lookahead MV, the
 *                              propagated amount, and its bilinear split over the (up to) four
 *                              importance blocks of the reference that the displaced block covers
 *   r1o_importance_finalize    :1054-1071 importance = log2(1 + importance / intra_cost)
 * f32 arithmetic in the reference's order (this file is built with -ffp-contract=off); the `+=` into
 * the reference's array happens in raster order of the source blocks, TL, TR, BL, BR within a block.
 * Parity UNPINNED: the reference holds no test for this function.
 */
#include <math.h>

#include "r1o.h"

#define IMP_B 8                      /* IMPORTANCE_BLOCK_SIZE, src/api/internal.rs:24 */
#define B_MV ((int64_t)IMP_B * 8)    /* BLOCK_SIZE_IN_MV_UNITS */
#define A_MV (B_MV * B_MV)           /* BLOCK_AREA_IN_MV_UNITS */

static void propagate(float *ref_imp, int w, int h, int64_t bx_mv, int64_t by_mv, float amount, float fraction) {
  const int64_t x = bx_mv / B_MV, y = by_mv / B_MV; /* i64 `/`: truncation (operands are multiples of B_MV) */
  if (x >= 0 && y >= 0 && x < w && y < h) ref_imp[y * w + x] += amount * fraction;
}

/* mvs: the lookahead MV field of this reference, MI (4x4) granular, mv_cols entries per row:
 * fi.lookahead_mvs[mv_index][y * 2][x * 2] = mvs[(2 y) * mv_cols + 2 x], as {row, col} i16 pairs */
void r1o_importance_propagate(const void *frame0, const r1o_plane_cfg *fcfg, const void *ref0, const r1o_plane_cfg *rcfg,
                              const int16_t *mvs, int mv_cols, const uint32_t *intra_costs, const float *importances,
                              int n_refs, int w, int h, int bit_depth, int px_bytes, float *ref_importances) {
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      const int16_t *mv = mvs + 2 * ((int64_t)(2 * y) * mv_cols + 2 * x);
      const int64_t reference_x = (int64_t)x * B_MV + mv[1], reference_y = (int64_t)y * B_MV + mv[0];
      const int64_t rx = reference_x / 8, ry = reference_y / 8; /* `as isize / 8`: toward zero */
      const uint8_t *po = (const uint8_t *)frame0 + ((int64_t)(y * IMP_B) * fcfg->stride + x * IMP_B) * px_bytes;
      const uint8_t *pr = (const uint8_t *)ref0 + (ry * rcfg->stride + rx) * px_bytes;
      const float inter_cost = (float)r1o_satd(po, fcfg->stride, pr, rcfg->stride, IMP_B, IMP_B, px_bytes);
      const float intra_cost = (float)intra_costs[y * w + x];
      const float future_importance = importances[y * w + x];
      float propagate_fraction = 1.f - inter_cost / intra_cost;
      propagate_fraction = fmaxf(propagate_fraction, 0.f); /* f32::max: the non-NaN operand */
      const float propagate_amount = (intra_cost + future_importance) * propagate_fraction / (float)n_refs;

      const int64_t tlx = (reference_x - (reference_x < 0 ? B_MV - 1 : 0)) / B_MV * B_MV;
      const int64_t tly = (reference_y - (reference_y < 0 ? B_MV - 1 : 0)) / B_MV * B_MV;
      const int64_t trx = tlx + B_MV, try_ = tly, blx = tlx, bly = tly + B_MV, brx = trx, bry = bly;
      propagate(ref_importances, w, h, tlx, tly, propagate_amount,
                (float)((trx - reference_x) * (bly - reference_y)) / (float)A_MV);
      propagate(ref_importances, w, h, trx, try_, propagate_amount,
                (float)((reference_x + B_MV - trx) * (bly - reference_y)) / (float)A_MV);
      propagate(ref_importances, w, h, blx, bly, propagate_amount,
                (float)((trx - reference_x) * (reference_y + B_MV - bly)) / (float)A_MV);
      propagate(ref_importances, w, h, brx, bry, propagate_amount,
                (float)((reference_x + B_MV - trx) * (reference_y + B_MV - bly)) / (float)A_MV);
    }
}

void r1o_importance_finalize(const uint32_t *intra_costs, float *importances, int64_t n) {
  for (int64_t i = 0; i < n; i++) {
    const float intra_cost = (float)intra_costs[i];
    importances[i] = intra_cost > 0.f ? log2f(1.f + importances[i] / intra_cost) : 0.f;
  }
}
