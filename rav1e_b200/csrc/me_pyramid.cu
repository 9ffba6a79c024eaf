// me_pyramid.cu -- the full-pel stages of FullSearch motion estimation on the frame pyramid,
// batched over blocks (host-side window arithmetic; the searches run in me_fs.cu):
//   estimate_motion_ss4        src/me.rs:1023-1075   quarter-res full search, +-192/64 px * me_range_scale
//   estimate_motion_ss2 + FullSearch::me_ss2   src/me.rs:280-333, 597-650   half-res, +-16 px around <= 3 pmvs
//   estimate_motion + FullSearch::full_pixel_me  src/me.rs:335-385, 522-565 full-res, +-16 px step 2 around <= 7 cmvs
// Window arithmetic follows the reference literally: `/ 8` truncates toward zero, `>> n` is an
// arithmetic shift, clamps in source order.  lambda = (me_lambda * 256 * k) as u32 with
// k = 1/16 * 0.125, 1/4 * 0.125, 0.5 (src/me.rs:1051, 302, 358).
#include <vector>

#include "common.cuh"

namespace {

// adjust_bo, src/me.rs:993-1003
inline void adjust_bo(int bo_x, int bo_y, int mi_w, int mi_h, int blk_w, int blk_h, int &ax, int &ay) {
  ax = bo_x < mi_w - blk_w / 4 ? bo_x : mi_w - blk_w / 4;
  if (ax < 0) ax = 0;
  ay = bo_y < mi_h - blk_h / 4 ? bo_y : mi_h - blk_h / 4;
  if (ay < 0) ay = 0;
}

// BlockSize::from_width_and_height; -1 when no such block size exists
inline int bsize_from_wh(int w, int h) {
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++)
    if (blk_w_of(b) == w && blk_h_of(b) == h) return b;
  return -1;
}

inline int64_t imax(int64_t a, int64_t b) { return a > b ? a : b; }
inline int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }

int check_common(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_block_job *jobs, int64_t n, r1c_mv *out) {
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 27) && (n == 0 || (jobs && out)), R1C_ERR_ARG, "bad job/result array");
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, jobs[i].bsize >= 0 && jobs[i].bsize < R1C_BLOCK_SIZES_ALL && jobs[i].bo_x >= 0 && jobs[i].bo_y >= 0, R1C_ERR_ARG,
                "block job %lld: bad block size or offset", (long long)i);
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_estimate_motion_ss4_batch(r1c_ctx *ctx, const r1c_plane *org_qres, const r1c_plane *ref_qres, int w_in_b,
                                             int h_in_b, const r1c_block_job *jobs, int64_t n, int me_range_scale,
                                             double me_lambda, int allow_high_precision_mv, int bit_depth, r1c_mv *out) {
  int rc = check_common(ctx, org_qres, ref_qres, jobs, n, out);
  if (rc || n == 0) return rc;
  const uint32_t lambda = (uint32_t)(me_lambda * 256.0 / 16.0 * 0.125);
  std::vector<r1c_fs_job> fs((size_t)n);
  std::vector<r1c_me_result> res((size_t)n);
  for (int64_t i = 0; i < n; i++) {
    const int bw = blk_w_of(jobs[i].bsize), bh = blk_h_of(jobs[i].bsize);
    const int qb = bsize_from_wh(bw >> 2, bh >> 2);
    R1C_REQUIRE(ctx, qb >= 0, R1C_ERR_ARG, "block job %lld: %dx%d has no quarter-resolution block size", (long long)i, bw, bh);
    int ax, ay;
    adjust_bo(jobs[i].bo_x, jobs[i].bo_y, w_in_b, h_in_b, bw, bh, ax, ay);
    const int64_t po_x = ((int64_t)ax << 2) >> 2, po_y = ((int64_t)ay << 2) >> 2;  // << BLOCK_TO_PLANE_SHIFT >> 2
    int64_t r[4];
    r1c_mv_range(w_in_b, h_in_b, ax, ay, bw, bh, r);
    const int64_t range_x = 192 * (int64_t)me_range_scale, range_y = 64 * (int64_t)me_range_scale;
    r1c_fs_job &f = fs[(size_t)i];
    f.po_x = (int32_t)po_x;
    f.po_y = (int32_t)po_y;
    f.bsize = qb;
    f.x_lo = (int32_t)(po_x + (imax(-range_x, r[0] / 8) >> 2));
    f.x_hi = (int32_t)(po_x + (imin(range_x, r[1] / 8) >> 2));
    f.y_lo = (int32_t)(po_y + (imax(-range_y, r[2] / 8) >> 2));
    f.y_hi = (int32_t)(po_y + (imin(range_y, r[3] / 8) >> 2));
    f.step = 1;
    f.lambda = lambda;
    f.pmv[0] = f.pmv[1] = r1c_mv{0, 0};
    f.allow_high_precision_mv = allow_high_precision_mv;
    f.has_init = 0;
    res[(size_t)i].mv = r1c_mv{0, 0};
    res[(size_t)i].reserved = 0;
    res[(size_t)i].cost = ~0ull;
  }
  if ((rc = r1c_full_search_batch(ctx, org_qres, ref_qres, fs.data(), n, bit_depth, res.data()))) return rc;
  for (int64_t i = 0; i < n; i++) out[i] = r1c_mv{(int16_t)(res[(size_t)i].mv.row * 4), (int16_t)(res[(size_t)i].mv.col * 4)};
  return R1C_OK;
}

// shared by ss2 (shift 1, step 1, half-res block) and the full-res stage (shift 0, step 2): one round per
// predictor slot; blocks without a predictor in that slot sit the round out
static int predictor_rounds(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int w_in_b, int h_in_b,
                            const r1c_block_job *jobs, int64_t n, const r1c_mv *preds, const uint8_t *valid, int slots, int shift,
                            int step, uint32_t lambda, int allow_hp, int bit_depth, std::vector<r1c_me_result> &best) {
  best.assign((size_t)n, r1c_me_result{r1c_mv{0, 0}, 0, ~0ull});
  std::vector<r1c_fs_job> fs;
  std::vector<r1c_me_result> res;
  std::vector<int64_t> idx;
  for (int s = 0; s < slots; s++) {
    fs.clear();
    res.clear();
    idx.clear();
    for (int64_t i = 0; i < n; i++) {
      if (!valid[i * slots + s]) continue;
      const int bw = blk_w_of(jobs[i].bsize), bh = blk_h_of(jobs[i].bsize);
      const int sb = bsize_from_wh(bw >> shift, bh >> shift);
      R1C_REQUIRE(ctx, sb >= 0, R1C_ERR_ARG, "block job %lld: %dx%d has no block size at this pyramid level", (long long)i, bw, bh);
      int ax, ay;
      adjust_bo(jobs[i].bo_x, jobs[i].bo_y, w_in_b, h_in_b, bw, bh, ax, ay);
      const int64_t po_x = ((int64_t)ax << 2) >> shift, po_y = ((int64_t)ay << 2) >> shift;
      int64_t r[4];
      r1c_mv_range(w_in_b, h_in_b, ax, ay, bw, bh, r);
      const r1c_mv p = preds[i * slots + s];
      const int64_t range = 16, pc = p.col / 8, pr = p.row / 8;  // i16 / 8 truncates toward zero
      r1c_fs_job f;
      f.po_x = (int32_t)po_x;
      f.po_y = (int32_t)po_y;
      f.bsize = sb;
      f.x_lo = (int32_t)(po_x + (imin(imax(pc - range, r[0] / 8), r[1] / 8) >> shift));
      f.x_hi = (int32_t)(po_x + (imin(imax(pc + range, r[0] / 8), r[1] / 8) >> shift));
      f.y_lo = (int32_t)(po_y + (imin(imax(pr - range, r[2] / 8), r[3] / 8) >> shift));
      f.y_hi = (int32_t)(po_y + (imin(imax(pr + range, r[2] / 8), r[3] / 8) >> shift));
      f.step = step;
      f.lambda = lambda;
      f.pmv[0] = f.pmv[1] = r1c_mv{0, 0};
      f.allow_high_precision_mv = allow_hp;
      f.has_init = 1;
      fs.push_back(f);
      res.push_back(best[(size_t)i]);
      idx.push_back(i);
    }
    if (fs.empty()) continue;
    const int rc = r1c_full_search_batch(ctx, org, ref, fs.data(), (int64_t)fs.size(), bit_depth, res.data());
    if (rc) return rc;
    for (size_t k = 0; k < idx.size(); k++) best[(size_t)idx[k]] = res[k];
  }
  return R1C_OK;
}

extern "C" int r1c_estimate_motion_ss2_batch(r1c_ctx *ctx, const r1c_plane *org_hres, const r1c_plane *ref_hres, int w_in_b,
                                             int h_in_b, const r1c_block_job *jobs, int64_t n, const r1c_mv *pmvs,
                                             const uint8_t *pmv_valid, double me_lambda, int allow_high_precision_mv,
                                             int bit_depth, r1c_mv *out) {
  int rc = check_common(ctx, org_hres, ref_hres, jobs, n, out);
  if (rc || n == 0) return rc;
  R1C_REQUIRE(ctx, pmvs && pmv_valid, R1C_ERR_ARG, "null predictor arrays");
  std::vector<r1c_me_result> best;
  const uint32_t lambda = (uint32_t)(me_lambda * 256.0 / 4.0 * 0.125);
  if ((rc = predictor_rounds(ctx, org_hres, ref_hres, w_in_b, h_in_b, jobs, n, pmvs, pmv_valid, 3, 1, 1, lambda,
                             allow_high_precision_mv, bit_depth, best)))
    return rc;
  for (int64_t i = 0; i < n; i++) out[i] = r1c_mv{(int16_t)(best[(size_t)i].mv.row * 2), (int16_t)(best[(size_t)i].mv.col * 2)};
  return R1C_OK;
}

extern "C" int r1c_estimate_motion_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int w_in_b, int h_in_b,
                                         const r1c_block_job *jobs, int64_t n, const r1c_mv *cmvs, const uint8_t *cmv_valid,
                                         double me_lambda, int allow_high_precision_mv, int bit_depth, r1c_mv *out) {
  int rc = check_common(ctx, org, ref, jobs, n, out);
  if (rc || n == 0) return rc;
  R1C_REQUIRE(ctx, cmvs && cmv_valid, R1C_ERR_ARG, "null predictor arrays");
  std::vector<r1c_me_result> best;
  const uint32_t lambda = (uint32_t)(me_lambda * 256.0 * 0.5);
  if ((rc = predictor_rounds(ctx, org, ref, w_in_b, h_in_b, jobs, n, cmvs, cmv_valid, 7, 0, 2, lambda, allow_high_precision_mv,
                             bit_depth, best)))
    return rc;
  for (int64_t i = 0; i < n; i++) out[i] = best[(size_t)i].mv;
  return R1C_OK;
}
