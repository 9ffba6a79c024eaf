// subpel.cu -- sub-pel motion-vector evaluation and the telescopic sub-pel search.
//
// Reference semantics:
//   get_mv_rd_cost          src/me.rs:787-838  (range check, predict_inter into a
//                                               block-sized plane, SAD|SATD, rate)
//   compute_mv_rd_cost      src/me.rs:840-856  (cost = 256*dist + rate*lambda)
//   predict_inter/get_params src/predict.rs:255-338 (MV -> offset + 1/16 frac, clamp;
//                                               put_8tap with fi.default_filter =
//                                               REGULAR for both axes, encoder.rs:705)
//   telescopic_subpel_search src/me.rs:858-941  and the SATD re-evaluation that
//                                               precedes it, src/me.rs:232-253
//
// Round-1 structure: a composition of the verified building blocks, all on the
// device and without host synchronisation -- (1) a kernel turns (job, candidate)
// pairs into put_8tap jobs, (2) the MC kernel of mc.cu writes every predicted
// block into a dense scratch buffer, (3) one warp per pair computes SAD / SATD
// against the org block and the rate term.  The telescopic search runs its 3 (4)
// steps as kernel sequences; the per-step pick is a tiny kernel.
#include <cub/device/device_scan.cuh>

#include "dist_dev.cuh"
#include "me_common.cuh"

namespace {

struct SubArgs {
  PlaneView org, ref;
  const r1c_subpel_job *jobs;
  long long n;
  int ncand;
  int ncand_alloc;             // candidates the dense buffer was laid out for (>= ncand)
  const r1c_mv *cand;          // n * ncand
  const long long *area_off;   // exclusive prefix sum of block areas, n + 1 entries
  r1c_mc_job *mcjobs;          // n * ncand
  const uint8_t *dense;        // predicted blocks
  unsigned long long *costs;   // n * ncand
  r1c_me_result *results;
  int step;
  int bit_depth;
};

__global__ void area_kernel(const r1c_subpel_job *jobs, long long n, long long *areas) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) areas[i] = (long long)blk_w_of(jobs[i].bsize) * blk_h_of(jobs[i].bsize);
  if (i == n) areas[i] = 0;
}

__device__ __forceinline__ bool mv_in_range(const r1c_subpel_job &j, r1c_mv mv) {
  return mv.col >= j.mvx_min && mv.col <= j.mvx_max && mv.row >= j.mvy_min && mv.row <= j.mvy_max;
}

// candidate -> put_8tap job (get_params, src/predict.rs:267-284, luma: xdec = ydec = 0)
__device__ __forceinline__ r1c_mc_job make_mc_job(const PlaneView &ref, const r1c_subpel_job &j, r1c_mv mv,
                                                  long long tmp_off, bool valid) {
  r1c_mc_job m;
  const int bw = blk_w_of(j.bsize), bh = blk_h_of(j.bsize);
  const int ro = (int)mv.row >> 3, co = (int)mv.col >> 3;  // arithmetic shift: floor
  const int rf = ((int)mv.row - (ro << 3)) << 1, cf = ((int)mv.col - (co << 3)) << 1;
  int qx = j.po_x + co - 3, qy = j.po_y + ro - 3;
  qx = max(min(qx, ref.width), -ref.xorigin);  // PlaneSlice::clamp, src/frame/plane.rs:521-533
  qy = max(min(qy, ref.height), -ref.yorigin);
  m.dst_x = 0;
  m.dst_y = 0;
  m.src_x = qx + 3;
  m.src_y = qy + 3;
  m.w = (int16_t)(valid ? bw : 0);  // w == 0: the MC kernel does nothing for this job
  m.h = (int16_t)(valid ? bh : 0);
  m.col_frac = (int8_t)cf;
  m.row_frac = (int8_t)rf;
  m.mode_x = R1C_REGULAR;
  m.mode_y = R1C_REGULAR;
  m.tmp_off = (int32_t)tmp_off;
  return m;
}

// the source footprint of the prediction must be inside the allocation
__device__ __forceinline__ bool mc_footprint_ok(const PlaneView &v, const r1c_mc_job &m, int bw, int bh) {
  const int x = m.src_x - (m.col_frac ? 3 : 0), y = m.src_y - (m.row_frac ? 3 : 0);
  const int w = bw + (m.col_frac ? 7 : 0), h = bh + (m.row_frac ? 7 : 0);
  return x >= -v.xorigin && y >= -v.yorigin && v.xorigin + x + w <= v.alloc_w && v.yorigin + y + h <= v.alloc_h;
}

// explicit candidate list (r1c_subpel_cost_batch)
__global__ void gen_jobs_kernel(SubArgs a) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= a.n * a.ncand) return;
  const long long i = p / a.ncand;
  const int k = (int)(p - i * a.ncand);
  const r1c_subpel_job j = a.jobs[i];
  const r1c_mv mv = a.cand[p];
  const long long area = a.area_off[i + 1] - a.area_off[i];
  r1c_mc_job m = make_mc_job(a.ref, j, mv, a.area_off[i] * a.ncand + k * area, true);
  const bool valid = mv_in_range(j, mv) && mc_footprint_ok(a.ref, m, blk_w_of(j.bsize), blk_h_of(j.bsize));
  if (!valid) {
    m.w = 0;
    m.h = 0;
  }
  a.mcjobs[p] = m;
}

// candidates of one telescopic step: the 8 neighbours of results[i].mv at distance
// `step`, (i, j) raster order without the centre (src/me.rs:883-902)
__global__ void gen_step_kernel(SubArgs a, r1c_mv *cand_out) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= a.n * 8) return;
  const long long i = p >> 3;
  const int k = (int)(p & 7);
  const int kk = k < 4 ? k : k + 1;  // skip (1,1)
  const int ii = kk / 3, jj = kk - ii * 3;
  const r1c_subpel_job j = a.jobs[i];
  const r1c_mv c0 = a.results[i].mv;
  r1c_mv mv;
  mv.row = (int16_t)(c0.row + a.step * (ii - 1));
  mv.col = (int16_t)(c0.col + a.step * (jj - 1));
  cand_out[p] = mv;
  const long long area = a.area_off[i + 1] - a.area_off[i];
  r1c_mc_job m = make_mc_job(a.ref, j, mv, a.area_off[i] * 8 + k * area, true);
  const bool valid = mv_in_range(j, mv) && mc_footprint_ok(a.ref, m, blk_w_of(j.bsize), blk_h_of(j.bsize));
  if (!valid) {
    m.w = 0;
    m.h = 0;
  }
  a.mcjobs[p] = m;
}

// one warp per (job, candidate): distortion of the dense predicted block + rate
template <int PB>
__global__ void __launch_bounds__(256) cost_kernel(SubArgs a) {
  const int lane = threadIdx.x & 31;
  const long long p = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= a.n * a.ncand) return;
  const long long i = p / a.ncand;
  const r1c_subpel_job j = a.jobs[i];
  const r1c_mc_job m = a.mcjobs[p];
  if (m.w == 0) {  // out of range (or footprint outside the allocation): u64::MAX, src/me.rs:794-799
    if (lane == 0) a.costs[p] = ~0ull;
    return;
  }
  const int bw = blk_w_of(j.bsize), bh = blk_h_of(j.bsize);
  const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.po_y * a.org.stride + j.po_x) * PB;
  const uint8_t *pp = a.dense + (long long)m.tmp_off * PB;
  const uint32_t d = j.use_satd ? distdev::satd_block_warp<PB>(po, a.org.stride * PB, pp, bw * PB, bw, bh, lane)
                                : distdev::sad_warp<PB>(po, a.org.stride * PB, pp, bw * PB, bw, bh, lane);
  if (lane == 0) {
    const r1c_mv mv = a.cand[p];
    const uint32_t rate = me::mv_rate2(mv, j.pmv[0], j.pmv[1], j.allow_high_precision_mv);
    a.costs[p] = 256ull * d + (unsigned long long)rate * j.lambda;
  }
}

// src/me.rs:232-253: with use_satd the cost of the full-pel best_mv is re-evaluated
// through the full-pel path of get_mv_rd_cost (:828-836): ref block at po + mv/8
template <int PB>
__global__ void __launch_bounds__(256) refine_init_kernel(SubArgs a) {
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= a.n) return;
  const r1c_subpel_job j = a.jobs[i];
  if (!j.use_satd) return;
  const r1c_mv mv = a.results[i].mv;
  unsigned long long cost = ~0ull;
  const int bw = blk_w_of(j.bsize), bh = blk_h_of(j.bsize);
  const int rx = j.po_x + mv.col / 8, ry = j.po_y + mv.row / 8;  // truncating division
  const bool inside = rx >= -a.ref.xorigin && ry >= -a.ref.yorigin && a.ref.xorigin + rx + bw <= a.ref.alloc_w &&
                      a.ref.yorigin + ry + bh <= a.ref.alloc_h;
  if (mv_in_range(j, mv) && inside) {
    const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.po_y * a.org.stride + j.po_x) * PB;
    const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)ry * a.ref.stride + rx) * PB;
    const uint32_t d = distdev::satd_block_warp<PB>(po, a.org.stride * PB, pr, a.ref.stride * PB, bw, bh, lane);
    const uint32_t rate = me::mv_rate2(mv, j.pmv[0], j.pmv[1], j.allow_high_precision_mv);
    cost = 256ull * d + (unsigned long long)rate * j.lambda;
  }
  if (lane == 0) a.results[i].cost = cost;
}

// `if cost < *lowest_cost { lowest_cost = cost; best_mv = cand_mv }` over the 8
// candidates in order (src/me.rs:934-937)
__global__ void pick_kernel(SubArgs a, const r1c_mv *cand) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  r1c_me_result r = a.results[i];
  for (int k = 0; k < 8; k++) {
    const unsigned long long c = a.costs[i * 8 + k];
    if (c < r.cost) {
      r.cost = c;
      r.mv = cand[i * 8 + k];
    }
  }
  a.results[i] = r;
}

int check_sub(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int bit_depth, int64_t n) {
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes, R1C_ERR_ARG, "org/ref pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, org->px_bytes);
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 27), R1C_ERR_ARG, "bad job count");
  return R1C_OK;
}

// scratch slots 5.. belong to this file (0-1: host wrappers, 2-4: mc.cu)
int prepare(r1c_ctx *ctx, SubArgs &a, int ncand) {
  int rc;
  void *d_area, *d_off, *d_tmp, *d_mc, *d_cost;
  const size_t ab = (size_t)(a.n + 1) * sizeof(long long);
  if ((rc = r1c_scratch_dev(ctx, 5, ab, &d_area))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 6, ab, &d_off))) return rc;
  area_kernel<<<(unsigned)((a.n + 1 + 255) / 256), 256, 0, ctx->stream>>>(a.jobs, a.n, (long long *)d_area);
  R1C_LAUNCH_CHECK(ctx);
  size_t tb = 0;
  R1C_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, (long long *)d_area, (long long *)d_off, (int)(a.n + 1), ctx->stream));
  if ((rc = r1c_scratch_dev(ctx, 7, tb + 16, &d_tmp))) return rc;
  R1C_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tb, (long long *)d_area, (long long *)d_off, (int)(a.n + 1), ctx->stream));
  ctx->launches += 2;
  if ((rc = r1c_scratch_dev(ctx, 8, (size_t)a.n * ncand * sizeof(r1c_mc_job), &d_mc))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 9, (size_t)a.n * ncand * sizeof(unsigned long long), &d_cost))) return rc;
  void *d_dense;
  // the total block area is only known on the device: read it back (8 bytes) to size the
  // dense prediction buffer.  This is the one host synchronisation of the sub-pel path.
  long long total_px = 0;
  R1C_CUDA(ctx, cudaMemcpyAsync(&total_px, (long long *)d_off + a.n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  R1C_REQUIRE(ctx, total_px * ncand < (1ll << 31), R1C_ERR_ARG, "sub-pel batch too large (%lld px x %d candidates): split it", total_px, ncand);
  if ((rc = r1c_scratch_dev(ctx, 10, (size_t)total_px * ncand * 2 + 64, &d_dense))) return rc;
  a.area_off = (const long long *)d_off;
  a.mcjobs = (r1c_mc_job *)d_mc;
  a.costs = (unsigned long long *)d_cost;
  a.dense = (const uint8_t *)d_dense;
  a.ncand = ncand;
  return R1C_OK;
}

int run_costs(r1c_ctx *ctx, const r1c_plane *ref, SubArgs &a) {
  ctx->prof_mc_slot = R1C_PROF_SUBPEL_MC;
  int rc = r1c_mc_put_dense_dev(ctx, (void *)a.dense, ref, a.mcjobs, a.n * a.ncand, a.bit_depth);
  ctx->prof_mc_slot = 0;
  if (rc) return rc;
  const unsigned grid = (unsigned)((a.n * a.ncand + 7) / 8);
  r1c_prof_begin(ctx, R1C_PROF_SUBPEL_COST);
  if (ref->px_bytes == 1) cost_kernel<1><<<grid, 256, 0, ctx->stream>>>(a);
  else cost_kernel<2><<<grid, 256, 0, ctx->stream>>>(a);
  r1c_prof_end(ctx, R1C_PROF_SUBPEL_COST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

int refine_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_subpel_job *d_jobs, int64_t n,
               int bit_depth, r1c_me_result *d_results, int any_hp) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  SubArgs a = {};
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.jobs = d_jobs;
  a.n = n;
  a.results = d_results;
  a.bit_depth = bit_depth;
  int rc = prepare(ctx, a, 8);
  if (rc) return rc;
  void *d_cand;
  if ((rc = r1c_scratch_dev(ctx, 11, (size_t)n * 8 * sizeof(r1c_mv), &d_cand))) return rc;
  a.cand = (const r1c_mv *)d_cand;
  const unsigned wgrid = (unsigned)((n + 7) / 8);
  if (ref->px_bytes == 1) refine_init_kernel<1><<<wgrid, 256, 0, ctx->stream>>>(a);
  else refine_init_kernel<2><<<wgrid, 256, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  // steps 8, 4, 2 and 1 when high-precision MVs are allowed (src/me.rs:868-871).  The
  // flag is per frame in the reference; jobs of one batch share it (checked by the
  // host wrapper; the device entry point takes the first job's word for it).
  const int steps[4] = {8, 4, 2, 1};
  for (int s = 0; s < (any_hp ? 4 : 3); s++) {
    a.step = steps[s];
    gen_step_kernel<<<(unsigned)((n * 8 + 255) / 256), 256, 0, ctx->stream>>>(a, (r1c_mv *)d_cand);
    R1C_LAUNCH_CHECK(ctx);
    if ((rc = run_costs(ctx, ref, a))) return rc;
    pick_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, (const r1c_mv *)d_cand);
    R1C_LAUNCH_CHECK(ctx);
  }
  return R1C_OK;
}

// ---- diamond_me_search (src/me.rs:693-785) -----------------------------------------
// State per job: results[i] = (center_mv, center_mv_cost), radius[i] (0 = finished).
// One iteration = candidate generation, the cost kernels above, and the pick rule;
// the host reads back one counter per iteration (jobs still searching).

// get_mv_rd_cost without a temp plane (src/me.rs:826-836): block at po + mv / 8, truncating
template <int PB>
__global__ void __launch_bounds__(256) fullpel_cost_kernel(SubArgs a, const int *radius) {
  const int lane = threadIdx.x & 31;
  const long long p = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= a.n * a.ncand) return;
  const long long i = p / a.ncand;
  unsigned long long cost = ~0ull;
  if (!radius || radius[i] != 0) {
    const r1c_subpel_job j = a.jobs[i];
    const r1c_mv mv = a.cand[p];
    const int bw = blk_w_of(j.bsize), bh = blk_h_of(j.bsize);
    const int rx = j.po_x + mv.col / 8, ry = j.po_y + mv.row / 8;
    const bool inside = rx >= -a.ref.xorigin && ry >= -a.ref.yorigin && a.ref.xorigin + rx + bw <= a.ref.alloc_w &&
                        a.ref.yorigin + ry + bh <= a.ref.alloc_h;
    if (mv_in_range(j, mv) && inside) {
      const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.po_y * a.org.stride + j.po_x) * PB;
      const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)ry * a.ref.stride + rx) * PB;
      const uint32_t d = j.use_satd ? distdev::satd_block_warp<PB>(po, a.org.stride * PB, pr, a.ref.stride * PB, bw, bh, lane)
                                    : distdev::sad_warp<PB>(po, a.org.stride * PB, pr, a.ref.stride * PB, bw, bh, lane);
      const uint32_t rate = me::mv_rate2(mv, j.pmv[0], j.pmv[1], j.allow_high_precision_mv);
      cost = 256ull * d + (unsigned long long)rate * j.lambda;
    }
  }
  if (lane == 0) a.costs[p] = cost;
}

// the 4 diamond candidates around the centre, pattern order (1,0),(0,1),(-1,0),(0,-1) in (row, col)
__global__ void gen_diamond_kernel(SubArgs a, r1c_mv *cand_out, const int *radius, int subpixel) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= a.n * 4) return;
  const long long i = p >> 2;
  const int k = (int)(p & 3);
  const int pr = k == 0 ? 1 : (k == 2 ? -1 : 0), pc = k == 1 ? 1 : (k == 3 ? -1 : 0);
  const int rad = radius[i];
  const r1c_subpel_job j = a.jobs[i];
  const r1c_mv c0 = a.results[i].mv;
  r1c_mv mv;
  mv.row = (int16_t)(c0.row + rad * pr);
  mv.col = (int16_t)(c0.col + rad * pc);
  cand_out[p] = mv;
  if (subpixel) {
    const long long area = a.area_off[i + 1] - a.area_off[i];
    r1c_mc_job m = make_mc_job(a.ref, j, mv, a.area_off[i] * a.ncand_alloc + k * area, true);
    const bool valid = rad != 0 && mv_in_range(j, mv) && mc_footprint_ok(a.ref, m, blk_w_of(j.bsize), blk_h_of(j.bsize));
    if (!valid) {
      m.w = 0;
      m.h = 0;
    }
    a.mcjobs[p] = m;
  }
}

// the predictor list as put_8tap jobs (sub-pel get_best_predictor)
__global__ void gen_pred_jobs_kernel(SubArgs a) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= a.n * a.ncand) return;
  const long long i = p / a.ncand;
  const int k = (int)(p - i * a.ncand);
  const r1c_subpel_job j = a.jobs[i];
  const r1c_mv mv = a.cand[p];
  const long long area = a.area_off[i + 1] - a.area_off[i];
  r1c_mc_job m = make_mc_job(a.ref, j, mv, a.area_off[i] * a.ncand_alloc + k * area, true);
  if (!(mv_in_range(j, mv) && mc_footprint_ok(a.ref, m, blk_w_of(j.bsize), blk_h_of(j.bsize)))) {
    m.w = 0;
    m.h = 0;
  }
  a.mcjobs[p] = m;
}

// get_best_predictor, src/me.rs:655-691: first strict minimum over the predictors
__global__ void best_predictor_kernel(SubArgs a, int *radius, int radius0) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  r1c_me_result r;
  r.mv.row = 0;
  r.mv.col = 0;
  r.reserved = 0;
  r.cost = ~0ull;
  for (int k = 0; k < a.ncand; k++) {
    const unsigned long long c = a.costs[i * a.ncand + k];
    if (c < r.cost) {
      r.cost = c;
      r.mv = a.cand[i * a.ncand + k];
    }
  }
  a.results[i] = r;
  radius[i] = radius0;
}

// src/me.rs:737-781: best of the 4 (strict <), then keep / shrink / move
__global__ void diamond_pick_kernel(SubArgs a, const r1c_mv *cand, int *radius, int radius_end, int *active) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  int rad = radius[i];
  if (rad == 0) return;
  unsigned long long best = ~0ull;
  r1c_mv best_mv;
  best_mv.row = 0;
  best_mv.col = 0;
  for (int k = 0; k < 4; k++) {
    const unsigned long long c = a.costs[i * 4 + k];
    if (c < best) {
      best = c;
      best_mv = cand[i * 4 + k];
    }
  }
  r1c_me_result r = a.results[i];
  if (r.cost <= best) {
    rad = rad == radius_end ? 0 : rad / 2;
    radius[i] = rad;
  } else {
    r.mv = best_mv;
    r.cost = best;
    a.results[i] = r;
  }
  if (rad != 0) atomicAdd(active, 1);
}

int run_fullpel_costs(r1c_ctx *ctx, const r1c_plane *ref, SubArgs &a, const int *radius) {
  const unsigned grid = (unsigned)((a.n * a.ncand + 7) / 8);
  r1c_prof_begin(ctx, R1C_PROF_SUBPEL_COST);
  if (ref->px_bytes == 1) fullpel_cost_kernel<1><<<grid, 256, 0, ctx->stream>>>(a, radius);
  else fullpel_cost_kernel<2><<<grid, 256, 0, ctx->stream>>>(a, radius);
  r1c_prof_end(ctx, R1C_PROF_SUBPEL_COST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// d_pred: n * n_pred predictors; d_results: out
int diamond_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_subpel_job *d_jobs, int64_t n,
                int subpixel, int allow_hp, const r1c_mv *d_pred, int n_pred, int bit_depth, r1c_me_result *d_results) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  SubArgs a = {};
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.jobs = d_jobs;
  a.n = n;
  a.results = d_results;
  a.bit_depth = bit_depth;
  const int nalloc = n_pred > 4 ? n_pred : 4;
  int rc = prepare(ctx, a, nalloc);
  if (rc) return rc;
  a.ncand_alloc = nalloc;
  void *d_cand, *d_state;
  if ((rc = r1c_scratch_dev(ctx, 11, (size_t)n * 4 * sizeof(r1c_mv), &d_cand))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 12, (size_t)(n + 1) * sizeof(int), &d_state))) return rc;
  int *radius = (int *)d_state, *active = (int *)d_state + n;
  // get_best_predictor
  a.cand = d_pred;
  a.ncand = n_pred;
  if (subpixel) {
    gen_pred_jobs_kernel<<<(unsigned)((n * n_pred + 255) / 256), 256, 0, ctx->stream>>>(a);
    R1C_LAUNCH_CHECK(ctx);
    if ((rc = run_costs(ctx, ref, a))) return rc;
  } else if ((rc = run_fullpel_costs(ctx, ref, a, nullptr))) {
    return rc;
  }
  const int radius0 = subpixel ? 4 : 16, radius_end = subpixel ? (allow_hp ? 1 : 2) : 8;
  best_predictor_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, radius, radius0);
  R1C_LAUNCH_CHECK(ctx);
  a.cand = (const r1c_mv *)d_cand;
  a.ncand = 4;
  int *h_active;
  if ((rc = r1c_stage_host(ctx, 3, sizeof(int), (void **)&h_active))) return rc;
  for (int it = 0;; it++) {
    R1C_REQUIRE(ctx, it < 100000, R1C_ERR_CUDA, "diamond search did not converge");
    R1C_CUDA(ctx, cudaMemsetAsync(active, 0, sizeof(int), ctx->stream));
    gen_diamond_kernel<<<(unsigned)((n * 4 + 255) / 256), 256, 0, ctx->stream>>>(a, (r1c_mv *)d_cand, radius, subpixel);
    R1C_LAUNCH_CHECK(ctx);
    if (subpixel) rc = run_costs(ctx, ref, a);
    else rc = run_fullpel_costs(ctx, ref, a, radius);
    if (rc) return rc;
    diamond_pick_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, (const r1c_mv *)d_cand, radius, radius_end, active);
    R1C_LAUNCH_CHECK(ctx);
    R1C_CUDA(ctx, cudaMemcpyAsync(h_active, active, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (*h_active == 0) break;
  }
  return R1C_OK;
}

int validate_jobs(r1c_ctx *ctx, const r1c_plane *org, const r1c_subpel_job *jobs, int64_t n, long long *total_px,
                  int *hp) {
  long long t = 0;
  for (int64_t i = 0; i < n; i++) {
    const r1c_subpel_job &j = jobs[i];
    R1C_REQUIRE(ctx, j.bsize >= 0 && j.bsize < R1C_BLOCK_SIZES_ALL, R1C_ERR_ARG, "subpel job %lld: bad block size", (long long)i);
    R1C_REQUIRE(ctx, region_in_plane(org->cfg, j.po_x, j.po_y, blk_w_of(j.bsize), blk_h_of(j.bsize)), R1C_ERR_BOUNDS,
                "subpel job %lld: org block leaves the allocated plane", (long long)i);
    R1C_REQUIRE(ctx, (j.allow_high_precision_mv != 0) == (jobs[0].allow_high_precision_mv != 0), R1C_ERR_ARG,
                "subpel jobs of one batch must share allow_high_precision_mv");
    t += (long long)blk_w_of(j.bsize) * blk_h_of(j.bsize);
  }
  *total_px = t;
  *hp = n > 0 && jobs[0].allow_high_precision_mv != 0;
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_subpel_cost_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                     const r1c_subpel_job *jobs, int64_t n, const r1c_mv *cand, int n_cand,
                                     int bit_depth, uint64_t *costs) {
  int rc = check_sub(ctx, org, ref, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n_cand >= 1 && n_cand <= 64 && (n == 0 || (jobs && cand && costs)), R1C_ERR_ARG, "bad candidate list");
  if (n == 0) return R1C_OK;
  long long total_px;
  int hp;
  if ((rc = validate_jobs(ctx, org, jobs, n, &total_px, &hp))) return rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *h_jobs, *h_cand, *d_jobs, *d_cand;
  const size_t jb = (size_t)n * sizeof(r1c_subpel_job), cb = (size_t)n * n_cand * sizeof(r1c_mv);
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, cb, &h_cand))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, cb, &d_cand))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_cand, cand, cb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_cand, h_cand, cb, cudaMemcpyHostToDevice, ctx->stream));
  SubArgs a = {};
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.jobs = (const r1c_subpel_job *)d_jobs;
  a.n = n;
  a.cand = (const r1c_mv *)d_cand;
  a.bit_depth = bit_depth;
  if ((rc = prepare(ctx, a, n_cand))) return rc;
  gen_jobs_kernel<<<(unsigned)((n * n_cand + 255) / 256), 256, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  if ((rc = run_costs(ctx, ref, a))) return rc;
  void *h_cost;
  const size_t ob = (size_t)n * n_cand * sizeof(uint64_t);
  if ((rc = r1c_stage_host(ctx, 2, ob, &h_cost))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_cost, a.costs, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(costs, h_cost, ob);
  return R1C_OK;
}

extern "C" int r1c_subpel_refine_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                           const r1c_subpel_job *d_jobs, int64_t n, int bit_depth,
                                           r1c_me_result *d_results) {
  int rc = check_sub(ctx, org, ref, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (d_jobs && d_results), R1C_ERR_ARG, "null job/result array");
  if (n == 0) return R1C_OK;
  int hp = 0;
  R1C_CUDA(ctx, cudaMemcpyAsync(&hp, &d_jobs[0].allow_high_precision_mv, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return refine_dev(ctx, org, ref, d_jobs, n, bit_depth, d_results, hp != 0);
}

extern "C" int r1c_subpel_refine_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                       const r1c_subpel_job *jobs, int64_t n, int bit_depth,
                                       r1c_me_result *results) {
  int rc = check_sub(ctx, org, ref, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (jobs && results), R1C_ERR_ARG, "null job/result array");
  if (n == 0) return R1C_OK;
  long long total_px;
  int hp;
  if ((rc = validate_jobs(ctx, org, jobs, n, &total_px, &hp))) return rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *h_jobs, *h_res, *d_jobs, *d_res;
  const size_t jb = (size_t)n * sizeof(r1c_subpel_job), rb = (size_t)n * sizeof(r1c_me_result);
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, rb, &h_res))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, rb, &d_res))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_res, results, rb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_res, h_res, rb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = refine_dev(ctx, org, ref, (const r1c_subpel_job *)d_jobs, n, bit_depth, (r1c_me_result *)d_res, hp))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_res, d_res, rb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(results, h_res, rb);
  return R1C_OK;
}

extern "C" int r1c_diamond_search_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                            const r1c_subpel_job *d_jobs, int64_t n, int subpixel,
                                            int allow_high_precision_mv, const r1c_mv *d_predictors, int n_pred,
                                            int bit_depth, r1c_me_result *d_results) {
  int rc = check_sub(ctx, org, ref, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n_pred >= 1 && n_pred <= 64 && (n == 0 || (d_jobs && d_predictors && d_results)), R1C_ERR_ARG, "bad predictor list");
  return diamond_dev(ctx, org, ref, d_jobs, n, subpixel != 0, allow_high_precision_mv != 0, d_predictors, n_pred, bit_depth, d_results);
}

extern "C" int r1c_diamond_search_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                        const r1c_subpel_job *jobs, int64_t n, int subpixel, const r1c_mv *predictors,
                                        int n_pred, int bit_depth, r1c_me_result *results) {
  int rc = check_sub(ctx, org, ref, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n_pred >= 1 && n_pred <= 64 && (n == 0 || (jobs && predictors && results)), R1C_ERR_ARG, "bad predictor list");
  if (n == 0) return R1C_OK;
  long long total_px;
  int hp;
  if ((rc = validate_jobs(ctx, org, jobs, n, &total_px, &hp))) return rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *h_jobs, *h_pred, *h_res, *d_jobs, *d_pred, *d_res;
  const size_t jb = (size_t)n * sizeof(r1c_subpel_job), pb = (size_t)n * n_pred * sizeof(r1c_mv), rb = (size_t)n * sizeof(r1c_me_result);
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, pb, &h_pred))) return rc;
  if ((rc = r1c_stage_host(ctx, 2, rb, &h_res))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, pb, &d_pred))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 13, rb, &d_res))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_pred, predictors, pb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_pred, h_pred, pb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = diamond_dev(ctx, org, ref, (const r1c_subpel_job *)d_jobs, n, subpixel != 0, hp, (const r1c_mv *)d_pred, n_pred, bit_depth,
                        (r1c_me_result *)d_res)))
    return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_res, d_res, rb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(results, h_res, rb);
  return R1C_OK;
}
