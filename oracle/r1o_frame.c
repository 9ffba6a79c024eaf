/*
 * r1o_frame.c -- ORACLE (test infrastructure, never shipped): the reference's
 * per-block motion estimation looped over every block of a frame.
 *   MotionEstimation::motion_estimation (FullSearch)   src/me.rs:193-271
 *   FullSearch::full_pixel_me                           src/me.rs:533-546
 *   FullSearch::sub_pixel_me -> telescopic_subpel_search src/me.rs:548-575,858-941
 *   get_mv_rd_cost (full-pel path, tmp_plane_opt = None) src/me.rs:828-836
 * One block at a time, exactly like the encoder calls it; OpenMP only spreads
 * the independent blocks over the host threads (bench.py cpu_baseline /
 * --impl reference, and the whole-frame parity tests).  No new arithmetic.
 * Parity: UNPINNED (the reference has no ME tests); literal restatement.
 */
#include "r1o.h"
#include <omp.h>

static int64_t tdiv8(int64_t v) { return v / 8; } /* Rust isize `/`: truncates toward zero (C99 too) */
static int64_t clamp64(int64_t v, int64_t lo, int64_t hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* result layout of include/rav1e_cuda.h r1c_frame_me: per BlockSize of the
 * mask, (width / bw) x (height / bh) whole blocks in raster order */
int64_t r1o_frame_me_layout(int width, int height, uint32_t bsize_mask, int64_t off[22], int32_t cols[22],
                            int32_t rows[22]) {
  int64_t o = 0;
  for (int b = 0; b < 22; b++) {
    int bw = r1o_block_width(b), bh = r1o_block_height(b);
    if (!((bsize_mask >> b) & 1u) || bw > 64 || bh > 64) {
      off[b] = -1;
      cols[b] = rows[b] = 0;
      continue;
    }
    off[b] = o;
    cols[b] = width / bw;
    rows[b] = height / bh;
    o += (int64_t)cols[b] * rows[b];
  }
  return o;
}

/* one block: src/me.rs:193-271 with pmv = {0,0}, cmv = 0 */
void r1o_block_me(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0, const r1o_plane_cfg *rcfg,
                  int width, int height, int bsize, int64_t po_x, int64_t po_y, int range_x, int range_y,
                  uint32_t lambda, int refine, int use_satd, int allow_hp, int bd, int pb, r1o_mv *out_mv,
                  uint64_t *out_cost) {
  const int bw = r1o_block_width(bsize), bh = r1o_block_height(bsize);
  int64_t r[4];
  r1o_mv_range(width / 4, height / 4, po_x / 4, po_y / 4, bw, bh, r);
  const r1o_mv pmv[2] = {{0, 0}, {0, 0}};
  /* src/me.rs:533-546 */
  const int64_t x_lo = po_x + clamp64(-range_x, tdiv8(r[0]), tdiv8(r[1]));
  const int64_t x_hi = po_x + clamp64(range_x, tdiv8(r[0]), tdiv8(r[1]));
  const int64_t y_lo = po_y + clamp64(-range_y, tdiv8(r[2]), tdiv8(r[3]));
  const int64_t y_hi = po_y + clamp64(range_y, tdiv8(r[2]), tdiv8(r[3]));
  r1o_mv best = {0, 0};
  uint64_t lowest = UINT64_MAX;
  r1o_full_search(org0, ocfg, ref0, rcfg, x_lo, x_hi, y_lo, y_hi, bw, bh, po_x, po_y, 1, lambda, pmv, allow_hp, pb,
                  &best, &lowest);
  if (refine) {
    if (use_satd) { /* src/me.rs:235-252: full-pel get_mv_rd_cost with SATD */
      if (best.col < r[0] || best.col > r[1] || best.row < r[2] || best.row > r[3]) {
        lowest = UINT64_MAX;
      } else {
        const uint8_t *o = (const uint8_t *)org0 + (po_y * ocfg->stride + po_x) * pb;
        const uint8_t *p =
            (const uint8_t *)ref0 + ((po_y + best.row / 8) * rcfg->stride + (po_x + best.col / 8)) * pb;
        uint32_t d = r1o_satd(o, ocfg->stride, p, rcfg->stride, bw, bh, pb);
        uint32_t r1 = r1o_mv_rate(best, pmv[0], allow_hp), r2 = r1o_mv_rate(best, pmv[1], allow_hp) + 1;
        lowest = 256ull * d + (uint64_t)(r1 < r2 ? r1 : r2) * lambda;
      }
    }
    /* telescopic_subpel_search, src/me.rs:858-941 */
    const int steps[4] = {8, 4, 2, 1};
    for (int s = 0; s < (allow_hp ? 4 : 3); s++) {
      const r1o_mv center = best;
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
          if (i == 1 && j == 1) continue;
          r1o_mv cand;
          cand.row = (int16_t)(center.row + steps[s] * (i - 1));
          cand.col = (int16_t)(center.col + steps[s] * (j - 1));
          uint64_t c = r1o_subpel_cost(org0, ocfg, ref0, rcfg, po_x, po_y, bw, bh, cand, pmv, lambda, use_satd,
                                       allow_hp, r, bd, pb);
          if (c < lowest) {
            lowest = c;
            best = cand;
          }
        }
    }
  }
  *out_mv = best;
  *out_cost = lowest;
}

/* every block (or, with idx != NULL, the n_idx frame-layout slots listed) */
int64_t r1o_frame_me(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0, const r1o_plane_cfg *rcfg,
                     int width, int height, int range_x, int range_y, uint32_t lambda, uint32_t bsize_mask,
                     int refine, int use_satd, int allow_hp, int bd, int pb, const int64_t *idx, int64_t n_idx,
                     int16_t *out_mv, uint64_t *out_cost) {
  int64_t off[22];
  int32_t cols[22], rows[22];
  const int64_t total = r1o_frame_me_layout(width, height, bsize_mask, off, cols, rows);
  const int64_t n = idx ? n_idx : total;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t k = 0; k < n; k++) {
    const int64_t slot = idx ? idx[k] : k;
    int b = 21;
    while (b >= 0 && (off[b] < 0 || off[b] > slot)) b--;
    if (b < 0 || slot >= total) continue;
    const int64_t local = slot - off[b];
    const int by = (int)(local / cols[b]), bx = (int)(local % cols[b]);
    r1o_mv mv;
    uint64_t cost;
    r1o_block_me(org0, ocfg, ref0, rcfg, width, height, b, (int64_t)bx * r1o_block_width(b),
                 (int64_t)by * r1o_block_height(b), range_x, range_y, lambda, refine, use_satd, allow_hp, bd, pb, &mv,
                 &cost);
    out_mv[2 * slot] = mv.row;
    out_mv[2 * slot + 1] = mv.col;
    out_cost[slot] = cost;
  }
  return n;
}
