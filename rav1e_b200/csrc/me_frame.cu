// me_frame.cu -- whole-frame motion estimation (r1c_frame_me*).
//
// Reference semantics: MotionEstimation::motion_estimation for FullSearch
// (src/me.rs:193-271) run for every whole block of every requested BlockSize
// of a frame against one reference frame, pmv = {0, 0}:
//   full_pixel_me   src/me.rs:533-546   window = clamp(+-range, mv range / 8)
//   full_search     src/me.rs:943-990   SAD cost, raster order, strict `<`
//   tail            src/me.rs:232-271   SATD re-evaluation + telescopic sub-pel
// It is a composition, entirely on the device, of the superblock search kernel
// (me_sb.cu: one pass over the window yields all 849 blocks of a 64x64 region)
// and of the sub-pel refinement (subpel.cu); this file only builds the job
// lists (once per frame geometry, kept resident) and moves results between the
// per-superblock layout and the per-BlockSize raster layout of the header.
#include <stdlib.h>

#include "me_common.cuh"

namespace {

struct Layout {
  int64_t off[R1C_BLOCK_SIZES_ALL];
  int32_t cols[R1C_BLOCK_SIZES_ALL], rows[R1C_BLOCK_SIZES_ALL];
  int64_t total;
};

bool params_ok(const r1c_frame_me_params *p) {
  return p && p->width >= 4 && p->height >= 4 && (p->width & 3) == 0 && (p->height & 3) == 0 &&
         p->width <= 16384 && p->height <= 16384 && p->range_x >= 0 && p->range_y >= 0 && p->bsize_mask != 0 &&
         (p->bsize_mask >> R1C_BLOCK_SIZES_ALL) == 0 && p->sb_row_lo >= 0 && p->sb_row_hi >= 0 &&
         (p->sb_row_hi == 0 || p->sb_row_hi > p->sb_row_lo) && p->sb_col_lo >= 0 && p->sb_col_hi >= 0 &&
         (p->sb_col_hi == 0 || p->sb_col_hi > p->sb_col_lo);
}

void make_layout(const r1c_frame_me_params *p, Layout *l) {
  int64_t o = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    const int bw = blk_w_of(b), bh = blk_h_of(b);
    if (!((p->bsize_mask >> b) & 1u) || bw > 64 || bh > 64) {
      l->off[b] = -1;
      l->cols[b] = l->rows[b] = 0;
      continue;
    }
    l->off[b] = o;
    l->cols[b] = p->width / bw;
    l->rows[b] = p->height / bh;
    o += (int64_t)l->cols[b] * l->rows[b];
  }
  l->total = o;
}

// results[map_dst[i]] = sb_results[map_src[i]]
__global__ void gather_kernel(const r1c_me_result *__restrict__ src, const int2 *__restrict__ map, long long n,
                              r1c_me_result *__restrict__ dst, int to_frame_layout) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int2 m = map[i];
  const uint4 v = *(const uint4 *)(src + m.x);
  *(uint4 *)(dst + (to_frame_layout ? (long long)m.y : i)) = v;
}
__global__ void gather_frame_kernel(const r1c_me_result *__restrict__ src, const int2 *__restrict__ map, long long n,
                                    r1c_me_result *__restrict__ dst) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  *(uint4 *)(dst + i) = *(const uint4 *)(src + map[i].y);
}
__global__ void scatter_kernel(const r1c_me_result *__restrict__ src, const int2 *__restrict__ map, long long n,
                               r1c_me_result *__restrict__ dst) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  *(uint4 *)(dst + map[i].y) = *(const uint4 *)(src + i);
}

bool is_whole(const r1c_frame_me_params *p) { return p->sb_row_hi == 0 && p->sb_col_hi == 0; }

void key_of(const r1c_frame_me_params *p, int32_t key[12]) {
  key[0] = p->width; key[1] = p->height; key[2] = p->range_x; key[3] = p->range_y;
  key[4] = (int32_t)p->lambda; key[5] = (int32_t)p->bsize_mask; key[6] = p->use_satd;
  key[7] = p->allow_high_precision_mv; key[8] = p->sb_row_lo; key[9] = p->sb_row_hi;
  key[10] = p->sb_col_lo; key[11] = p->sb_col_hi;
}

// job lists of one (frame geometry, parameters): superblock jobs, the
// (superblock slot -> frame slot) map and the sub-pel jobs of every block
int build_lists(r1c_ctx *ctx, const r1c_frame_me_params *p) {
  int32_t key[12];
  key_of(p, key);
  if (ctx->frame_me.valid && memcmp(key, ctx->frame_me.key, sizeof key) == 0) return R1C_OK;
  ctx->frame_me.valid = 0;
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (ctx->frame_me.d_sbjobs) cudaFree(ctx->frame_me.d_sbjobs);
  if (ctx->frame_me.d_map) cudaFree(ctx->frame_me.d_map);
  if (ctx->frame_me.d_subjobs) cudaFree(ctx->frame_me.d_subjobs);
  ctx->frame_me.d_sbjobs = ctx->frame_me.d_map = ctx->frame_me.d_subjobs = nullptr;

  Layout l;
  make_layout(p, &l);
  const int sbc = (p->width + 63) / 64, sbr = (p->height + 63) / 64;
  const int r_lo = p->sb_row_hi ? p->sb_row_lo : 0, r_hi = p->sb_row_hi ? (p->sb_row_hi < sbr ? p->sb_row_hi : sbr) : sbr;
  const int c_lo = p->sb_col_hi ? p->sb_col_lo : 0, c_hi = p->sb_col_hi ? (p->sb_col_hi < sbc ? p->sb_col_hi : sbc) : sbc;
  const int tc = c_hi > c_lo ? c_hi - c_lo : 0;  // superblock columns of the tile
  const int64_t n_sb = r_hi > r_lo ? (int64_t)(r_hi - r_lo) * tc : 0;
  r1c_sb_job *sb = (r1c_sb_job *)calloc((size_t)(n_sb ? n_sb : 1), sizeof(r1c_sb_job));
  // upper bound on the blocks of the shard
  int64_t cap = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++)
    if (l.off[b] >= 0 && n_sb) cap += (((int64_t)tc * 64) / blk_w_of(b) + 1) * (((int64_t)(r_hi - r_lo) * 64) / blk_h_of(b) + 1);
  int2 *map = (int2 *)malloc((size_t)(cap ? cap : 1) * sizeof(int2));
  r1c_subpel_job *sj = (r1c_subpel_job *)calloc((size_t)(cap ? cap : 1), sizeof(r1c_subpel_job));
  if (!sb || !map || !sj) {
    free(sb); free(map); free(sj);
    return r1c_set_err(ctx, R1C_ERR_NOMEM, "frame_me: host job lists");
  }
  for (int64_t i = 0; i < n_sb; i++) {
    sb[i].po_x = (int32_t)(c_lo + i % tc) * 64;
    sb[i].po_y = (int32_t)(r_lo + i / tc) * 64;
    sb[i].bsize_mask = p->bsize_mask;
  }
  const int w_in_b = p->width / 4, h_in_b = p->height / 4;
  int64_t n = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    if (l.off[b] < 0) continue;
    const int bw = blk_w_of(b), bh = blk_h_of(b);
    const int sb_off = r1c_sb_result_offset(b), per_row = 64 / bw;
    for (int by = 0; by < l.rows[b]; by++) {
      const int y = by * bh, sy = y / 64;
      if (sy < r_lo || sy >= r_hi) continue;
      for (int bx = 0; bx < l.cols[b]; bx++) {
        const int x = bx * bw, sx = x / 64;
        if (sx < c_lo || sx >= c_hi) continue;
        const int64_t sbi = (int64_t)(sy - r_lo) * tc + (sx - c_lo);
        map[n].x = (int)(sbi * R1C_SB_RESULTS + sb_off + ((y & 63) / bh) * per_row + (x & 63) / bw);
        map[n].y = (int)(l.off[b] + (int64_t)by * l.cols[b] + bx);
        r1c_subpel_job &j = sj[n];
        j.po_x = x;
        j.po_y = y;
        j.bsize = b;
        j.lambda = p->lambda;
        int64_t range[4];
        r1c_mv_range(w_in_b, h_in_b, x / 4, y / 4, bw, bh, range);
        j.mvx_min = (int32_t)range[0];
        j.mvx_max = (int32_t)range[1];
        j.mvy_min = (int32_t)range[2];
        j.mvy_max = (int32_t)range[3];
        j.use_satd = p->use_satd;
        j.allow_high_precision_mv = p->allow_high_precision_mv;
        n++;
      }
    }
  }
  int rc = R1C_OK;
  cudaError_t e = cudaSuccess;
  if (n_sb && e == cudaSuccess) e = cudaMalloc(&ctx->frame_me.d_sbjobs, (size_t)n_sb * sizeof(r1c_sb_job));
  if (n && e == cudaSuccess) e = cudaMalloc(&ctx->frame_me.d_map, (size_t)n * sizeof(int2));
  if (n && e == cudaSuccess) e = cudaMalloc(&ctx->frame_me.d_subjobs, (size_t)n * sizeof(r1c_subpel_job));
  if (n_sb && e == cudaSuccess) e = cudaMemcpy(ctx->frame_me.d_sbjobs, sb, (size_t)n_sb * sizeof(r1c_sb_job), cudaMemcpyHostToDevice);
  if (n && e == cudaSuccess) e = cudaMemcpy(ctx->frame_me.d_map, map, (size_t)n * sizeof(int2), cudaMemcpyHostToDevice);
  if (n && e == cudaSuccess) e = cudaMemcpy(ctx->frame_me.d_subjobs, sj, (size_t)n * sizeof(r1c_subpel_job), cudaMemcpyHostToDevice);
  free(sb); free(map); free(sj);
  if (e != cudaSuccess) rc = r1c_set_err(ctx, R1C_ERR_CUDA, "frame_me job lists: %s", cudaGetErrorString(e));
  if (rc) return rc;
  ctx->frame_me.n_sb = n_sb;
  ctx->frame_me.n_blocks = n;
  memcpy(ctx->frame_me.key, key, sizeof key);
  ctx->frame_me.valid = 1;
  return R1C_OK;
}

}  // namespace

extern "C" int64_t r1c_frame_me_count(const r1c_frame_me_params *p) {
  if (!params_ok(p)) return -1;
  Layout l;
  make_layout(p, &l);
  return l.total;
}

extern "C" int64_t r1c_frame_me_offset(const r1c_frame_me_params *p, int bsize, int32_t *cols, int32_t *rows) {
  if (!params_ok(p) || bsize < 0 || bsize >= R1C_BLOCK_SIZES_ALL) return -1;
  Layout l;
  make_layout(p, &l);
  if (cols) *cols = l.cols[bsize];
  if (rows) *rows = l.rows[bsize];
  return l.off[bsize];
}

extern "C" int r1c_frame_me_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                const r1c_frame_me_params *p, int bit_depth, r1c_me_result *d_results) {
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, params_ok(p), R1C_ERR_ARG, "frame_me: bad parameters");
  R1C_REQUIRE(ctx, d_results != nullptr, R1C_ERR_ARG, "frame_me: null result array");
  R1C_REQUIRE(ctx, org->cfg.width == p->width && org->cfg.height == p->height && ref->cfg.width == p->width &&
                       ref->cfg.height == p->height,
              R1C_ERR_ARG, "frame_me: plane size differs from the frame size");
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++)
    R1C_REQUIRE(ctx, !((p->bsize_mask >> b) & 1u) || (blk_w_of(b) <= 64 && blk_h_of(b) <= 64), R1C_ERR_ARG,
                "frame_me: block sizes above 64x64 are not searched as a whole");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = build_lists(ctx, p);
  if (rc) return rc;
  const int64_t n_sb = ctx->frame_me.n_sb, n = ctx->frame_me.n_blocks;
  if (n_sb == 0 || n == 0) return R1C_OK;
  r1c_sb_params sp;
  sp.w_in_b = p->width / 4;
  sp.h_in_b = p->height / 4;
  sp.range_x = p->range_x;
  sp.range_y = p->range_y;
  sp.step = 1;
  sp.lambda = p->lambda;
  sp.allow_high_precision_mv = p->allow_high_precision_mv;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes && bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG,
              "frame_me: bit_depth %d vs %d-byte pixels", bit_depth, org->px_bytes);
  const bool whole = is_whole(p);
  const unsigned grid = (unsigned)((n + 255) / 256);
  const r1c_sb_job *d_sbs = (const r1c_sb_job *)ctx->frame_me.d_sbjobs;
  const int2 *d_map = (const int2 *)ctx->frame_me.d_map;
  const r1c_subpel_job *d_sub = (const r1c_subpel_job *)ctx->frame_me.d_subjobs;
  // 1. full-pel search, scattered straight into the frame layout when me_sb2 takes it
  int direct = 0;
  if ((rc = r1c_sb2_search(ctx, org, ref, &sp, bit_depth, d_sbs, n_sb, d_results, d_map, n, &direct))) return rc;
  if (!direct) {
    void *d_sbres;
    if ((rc = r1c_scratch_dev(ctx, 12, (size_t)n_sb * R1C_SB_RESULTS * sizeof(r1c_me_result), &d_sbres))) return rc;
    if ((rc = r1c_me_sb_batch_dev(ctx, org, ref, &sp, d_sbs, n_sb, bit_depth, (r1c_me_result *)d_sbres))) return rc;
    gather_kernel<<<grid, 256, 0, ctx->stream>>>((const r1c_me_result *)d_sbres, d_map, n, d_results, 1);
    R1C_LAUNCH_CHECK(ctx);
  }
  if (!p->refine) return R1C_OK;
  // 2. SATD re-evaluation + telescopic sub-pel: fused kernel (8-bit), else the generic path
  {
    Layout l;
    make_layout(p, &l);
    long long off[R1C_BLOCK_SIZES_ALL];
    for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) off[b] = l.off[b];
    int fused = 0;
    if ((rc = r1c_refine_frame(ctx, org, ref, p, bit_depth, off, l.cols, l.rows, d_sbs, n_sb, d_results, &fused))) return rc;
    if (fused) return R1C_OK;
  }
  if (whole)  // job i <-> frame slot i: refine in place
    return r1c_subpel_refine_batch_dev(ctx, org, ref, d_sub, n, bit_depth, d_results);
  void *d_compact;
  if ((rc = r1c_scratch_dev(ctx, 13, (size_t)n * sizeof(r1c_me_result), &d_compact))) return rc;
  gather_frame_kernel<<<grid, 256, 0, ctx->stream>>>(d_results, d_map, n, (r1c_me_result *)d_compact);
  R1C_LAUNCH_CHECK(ctx);
  if ((rc = r1c_subpel_refine_batch_dev(ctx, org, ref, d_sub, n, bit_depth, (r1c_me_result *)d_compact))) return rc;
  scatter_kernel<<<grid, 256, 0, ctx->stream>>>((const r1c_me_result *)d_compact, d_map, n, d_results);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int64_t r1c_frame_me_shard_count(r1c_ctx *ctx, const r1c_frame_me_params *p) {
  if (!ctx || !params_ok(p)) return -1;
  if (cudaSetDevice(ctx->device) != cudaSuccess || build_lists(ctx, p)) return -1;
  return ctx->frame_me.n_blocks;
}

extern "C" int r1c_frame_me_compact_dev(r1c_ctx *ctx, const r1c_frame_me_params *p, const r1c_me_result *d_results,
                                        r1c_me_result *d_compact) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, params_ok(p), R1C_ERR_ARG, "frame_me: bad parameters");
  R1C_REQUIRE(ctx, d_results && d_compact, R1C_ERR_ARG, "frame_me: null result array");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = build_lists(ctx, p);
  if (rc) return rc;
  const int64_t n = ctx->frame_me.n_blocks;
  if (n == 0) return R1C_OK;
  gather_frame_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_results, (const int2 *)ctx->frame_me.d_map, n, d_compact);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// host result array: the shard's entries are computed on the device and copied back on the context stream.  Returns
// without synchronising when `results` is page-locked (*pending = 1: the copy is still in flight); pageable memory
// goes through the pinned stage and is complete on return.
static int frame_me_host(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_frame_me_params *p, int bit_depth,
                         r1c_me_result *results, int *pending) {
  *pending = 0;
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, params_ok(p), R1C_ERR_ARG, "frame_me: bad parameters");
  R1C_REQUIRE(ctx, results != nullptr, R1C_ERR_ARG, "frame_me: null result array");
  Layout l;
  make_layout(p, &l);
  const size_t rb = (size_t)l.total * sizeof(r1c_me_result);
  void *h_res = results, *d_res;
  int rc;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  // a page-locked result array is the DMA target itself; pageable memory goes through the pinned stage
  cudaPointerAttributes attr;
  const bool direct = cudaPointerGetAttributes(&attr, results) == cudaSuccess && attr.type == cudaMemoryTypeHost;
  if (!direct) {
    (void)cudaGetLastError();
    if ((rc = r1c_stage_host(ctx, 3, rb, &h_res))) return rc;
  }
  if ((rc = r1c_scratch_dev(ctx, 14, rb, &d_res))) return rc;
  const bool whole = is_whole(p);
  if (!whole) {  // entries outside the shard keep the caller's values
    if (!direct) memcpy(h_res, results, rb);
    R1C_CUDA(ctx, cudaMemcpyAsync(d_res, h_res, rb, cudaMemcpyHostToDevice, ctx->stream));
  }
  if ((rc = r1c_frame_me_dev(ctx, org, ref, p, bit_depth, (r1c_me_result *)d_res))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_res, d_res, rb, cudaMemcpyDeviceToHost, ctx->stream));
  if (direct) {
    *pending = 1;
    return R1C_OK;
  }
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(results, h_res, rb);
  return R1C_OK;
}

extern "C" int r1c_frame_me(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                            const r1c_frame_me_params *p, int bit_depth, r1c_me_result *results) {
  int pending = 0;
  const int rc = frame_me_host(ctx, org, ref, p, bit_depth, results, &pending);
  if (rc) return rc;
  if (pending) R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

// ---- a lookahead window of frame pairs in one call ----------------------------------------------------------
// The pairs go to two streams in turn (this context's and its peer's): the tail of one frame's kernels is filled
// by the next frame's, and with host buffers the copies of one frame overlap the kernels of another.
static int window_fork(r1c_ctx *ctx) {
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!ctx->peer) {
    ctx->peer = r1c_create(ctx->device);
    if (!ctx->peer) return r1c_set_err(ctx, R1C_ERR_NOMEM, "frame_me_window: cannot create the second stream");
  }
  if (!ctx->win_ev_created) {
    R1C_CUDA(ctx, cudaEventCreateWithFlags(&ctx->win_ev[0], cudaEventDisableTiming));
    R1C_CUDA(ctx, cudaEventCreateWithFlags(&ctx->win_ev[1], cudaEventDisableTiming));
    ctx->win_ev_created = 1;
  }
  // the peer starts after everything already queued on the context stream (the planes' uploads, ...)
  R1C_CUDA(ctx, cudaEventRecord(ctx->win_ev[0], ctx->stream));
  R1C_CUDA(ctx, cudaStreamWaitEvent(ctx->peer->stream, ctx->win_ev[0], 0));
  return R1C_OK;
}
static int window_join(r1c_ctx *ctx) {  // later work on the context stream sees the peer's results
  R1C_CUDA(ctx, cudaEventRecord(ctx->win_ev[1], ctx->peer->stream));
  R1C_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->win_ev[1], 0));
  return R1C_OK;
}
static int peer_error(r1c_ctx *ctx, int rc) {
  snprintf(ctx->err, sizeof ctx->err, "%s", ctx->peer->err);
  return rc;
}

extern "C" int r1c_frame_me_window_dev(r1c_ctx *ctx, const r1c_frame_pair *pairs, int64_t n, const r1c_frame_me_params *p,
                                       int bit_depth) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || pairs), R1C_ERR_ARG, "frame_me_window: null pair array");
  if (n == 0) return R1C_OK;
  int rc = window_fork(ctx);
  if (rc) return rc;
  for (int64_t i = 0; i < n; i++) {
    r1c_ctx *c = (i & 1) ? ctx->peer : ctx;
    if ((rc = r1c_frame_me_dev(c, pairs[i].org, pairs[i].ref, p, bit_depth, (r1c_me_result *)pairs[i].results)))
      return c == ctx ? rc : peer_error(ctx, rc);
  }
  return window_join(ctx);
}

extern "C" int r1c_frame_me_window(r1c_ctx *ctx, const r1c_frame_pair *pairs, int64_t n, const r1c_frame_me_params *p,
                                   int bit_depth) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || pairs), R1C_ERR_ARG, "frame_me_window: null pair array");
  if (n == 0) return R1C_OK;
  int rc = window_fork(ctx);
  if (rc) return rc;
  r1c_ctx *bad = nullptr;
  for (int64_t i = 0; i < n && !rc; i++) {
    r1c_ctx *c = (i & 1) ? ctx->peer : ctx;
    int pending;
    if (pairs[i].org_host) rc = r1c_plane_upload_async(c, (r1c_plane *)pairs[i].org, pairs[i].org_host);
    if (!rc && pairs[i].ref_host) rc = r1c_plane_upload_async(c, (r1c_plane *)pairs[i].ref, pairs[i].ref_host);
    if (!rc) rc = frame_me_host(c, pairs[i].org, pairs[i].ref, p, bit_depth, (r1c_me_result *)pairs[i].results, &pending);
    if (rc) bad = c;
  }
  if (rc) {  // drain what was queued before reporting
    cudaStreamSynchronize(ctx->peer->stream);
    cudaStreamSynchronize(ctx->stream);
    return bad == ctx->peer ? peer_error(ctx, rc) : rc;
  }
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->peer->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}
