/*
 * r1o_batch.c -- ORACLE (test infrastructure, never shipped): OpenMP loops of
 * the per-block oracle functions over whole job lists.  Used by bench.py's
 * `cpu_baseline` / `--impl reference` legs (all host threads) and by the
 * parity tests to check thousands of blocks quickly.  No new arithmetic here.
 */
#include "r1o.h"
#include <omp.h>

int r1o_num_threads(void) { return omp_get_max_threads(); }
/* launchers such as torchrun export OMP_NUM_THREADS=1; the timed CPU baseline asks for the cores explicitly */
void r1o_set_num_threads(int n) { omp_set_num_threads(n > 0 ? n : omp_get_num_procs()); }

/* jobs: n x {org_x, org_y, ref_x, ref_y, bsize} int32 */
void r1o_dist_batch(int use_satd, const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                    const r1o_plane_cfg *rcfg, const int32_t *jobs, int64_t n, int pb,
                    uint32_t *out) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < n; i++) {
    const int32_t *j = jobs + 5 * i;
    int w = r1o_block_width(j[4]), h = r1o_block_height(j[4]);
    const uint8_t *o = (const uint8_t *)org0 + ((int64_t)j[1] * ocfg->stride + j[0]) * pb;
    const uint8_t *r = (const uint8_t *)ref0 + ((int64_t)j[3] * rcfg->stride + j[2]) * pb;
    out[i] = use_satd ? r1o_satd(o, ocfg->stride, r, rcfg->stride, w, h, pb)
                      : r1o_sad(o, ocfg->stride, r, rcfg->stride, w, h, pb);
  }
}

/* jobs: n x {po_x, po_y, blk_w, blk_h, x_lo, x_hi, y_lo, y_hi, step, lambda,
 *            pmv0_row, pmv0_col, pmv1_row, pmv1_col, allow_hp, init} int32
 * (init: 0 = start from (MAX, mv 0); 1 = out[] already holds best/cost in)
 * out_mv: n x {row, col} int16 ; out_cost: n x u64 ; returns #failed jobs */
int r1o_full_search_batch(const void *org0, const r1o_plane_cfg *ocfg, const void *ref0,
                          const r1o_plane_cfg *rcfg, const int32_t *jobs, int64_t n, int pb,
                          int16_t *out_mv, uint64_t *out_cost) {
  int bad = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : bad)
  for (int64_t i = 0; i < n; i++) {
    const int32_t *j = jobs + 16 * i;
    r1o_mv pmv[2] = {{(int16_t)j[10], (int16_t)j[11]}, {(int16_t)j[12], (int16_t)j[13]}};
    r1o_mv best = {0, 0};
    uint64_t cost = UINT64_MAX;
    if (j[15]) { best.row = out_mv[2 * i]; best.col = out_mv[2 * i + 1]; cost = out_cost[i]; }
    int rc = r1o_full_search(org0, ocfg, ref0, rcfg, j[4], j[5], j[6], j[7], j[2], j[3], j[0], j[1],
                             j[8], (uint32_t)j[9], pmv, j[14], pb, &best, &cost);
    if (rc) bad++;
    out_mv[2 * i] = best.row;
    out_mv[2 * i + 1] = best.col;
    out_cost[i] = cost;
  }
  return bad;
}

/* jobs: n x {dst_x, dst_y, src_x, src_y, w, h, col_frac, row_frac, mode_x, mode_y} int32;
 * kind 0 put (dst plane), 1 prep (tmp + tmp_off[i]) */
void r1o_mc_batch(int kind, void *dst0, const r1o_plane_cfg *dcfg, int16_t *tmp,
                  const int64_t *tmp_off, const void *src0, const r1o_plane_cfg *scfg,
                  const int32_t *jobs, int64_t n, int bd, int pb) {
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = 0; i < n; i++) {
    const int32_t *j = jobs + 10 * i;
    const uint8_t *s = (const uint8_t *)src0 + ((int64_t)j[3] * scfg->stride + j[2]) * pb;
    if (kind == 0) {
      uint8_t *d = (uint8_t *)dst0 + ((int64_t)j[1] * dcfg->stride + j[0]) * pb;
      r1o_put_8tap(d, dcfg->stride, s, scfg->stride, j[4], j[5], j[6], j[7], j[8], j[9], bd, pb);
    } else {
      r1o_prep_8tap(tmp + tmp_off[i], s, scfg->stride, j[4], j[5], j[6], j[7], j[8], j[9], bd, pb);
    }
  }
}

/* jobs: n x {dst_x, dst_y, w, h} ; tmp1/tmp2 dense at off[i] */
void r1o_avg_batch(void *dst0, const r1o_plane_cfg *dcfg, const int16_t *t1, const int16_t *t2,
                   const int64_t *off, const int32_t *jobs, int64_t n, int bd, int pb) {
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = 0; i < n; i++) {
    const int32_t *j = jobs + 4 * i;
    uint8_t *d = (uint8_t *)dst0 + ((int64_t)j[1] * dcfg->stride + j[0]) * pb;
    r1o_mc_avg(d, dcfg->stride, t1 + off[i], t2 + off[i], j[2], j[3], bd, pb);
  }
}

/* n blocks of one (tx_size, tx_type), dense back to back */
int r1o_fht_batch(const int16_t *residual, int32_t *coeffs, int tx_size, int tx_type, int bd,
                  int64_t n) {
  if (!r1o_fht_valid(tx_size, tx_type)) return -1;
  int64_t area = (int64_t)r1o_tx_width(tx_size) * r1o_tx_height(tx_size);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++)
    r1o_fht(residual + i * area, coeffs + i * area, tx_size, tx_type, bd);
  return 0;
}

/* benches/dist.rs:101-118 (criterion's b.iter over one get_sad / get_satd call at region (0, 0)): `iters` back-to-back
 * calls on one thread, nanoseconds per call.  The result of every call is kept live through a volatile sink. */
#include <time.h>
double r1o_bench_dist(int use_satd, const void *org, ptrdiff_t os, const void *ref, ptrdiff_t rs, int w, int h, int pb,
                      int64_t iters) {
  volatile uint32_t sink = 0;
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int64_t i = 0; i < iters; i++) {
    const void *o = org;
    __asm__ volatile("" : "+r"(o));  /* the compiler may not hoist the call out of the loop */
    sink = use_satd ? r1o_satd(o, os, ref, rs, w, h, pb) : r1o_sad(o, os, ref, rs, w, h, pb);
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  (void)sink;
  return ((double)(t1.tv_sec - t0.tv_sec) * 1e9 + (double)(t1.tv_nsec - t0.tv_nsec)) / (double)iters;
}
