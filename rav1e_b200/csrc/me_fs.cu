// me_fs.cu -- generic batched full search: one job = one call of the
// reference's full_search (src/me.rs:943-990) for a block of any BlockSize, any
// window and any step.  Used for the quarter/half-resolution pyramid levels
// (estimate_motion_ss4, me_ss2) and as the general drop-in; the whole-superblock
// kernel in me_sb.cu is the fast path for dense full-resolution search.
//
// One CTA per job, one candidate position per thread at a time: the block of
// the `org` plane is staged in shared memory once (every thread reads the same
// word -> broadcast), the reference window is read through L1 with aligned
// 32-bit loads + funnel shift (neighbouring threads = neighbouring candidates,
// so the loads of a warp coalesce).  The argmin keeps the reference's order:
// raster scan, strict `<` (first minimum wins) == lexicographic min over
// (cost, raster index).
#include "me_common.cuh"

namespace {

constexpr int FS_THREADS = 256;

struct FsArgs {
  PlaneView org, ref;
  const r1c_fs_job *jobs;
  r1c_me_result *results;
  long long n;
};

struct Best {
  unsigned long long cost;
  uint32_t idx;
};
__device__ __forceinline__ bool better(unsigned long long c, uint32_t i, const Best &b) {
  return c < b.cost || (c == b.cost && i < b.idx);
}

template <int PB>
__global__ void __launch_bounds__(FS_THREADS) full_search_kernel(FsArgs a) {
  __shared__ uint32_t s_org[128 * 128 * 2 / 4];
  __shared__ Best s_best[FS_THREADS / 32];
  const long long ji = blockIdx.x;
  const r1c_fs_job j = a.jobs[ji];
  const int tid = threadIdx.x;
  const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
  const int step = j.step;

  // PlaneRegion::new asserts (src/tiling/plane_region.rs:139-142) for the search
  // region and the org block; a violating job is flagged instead of panicking
  bool ok = step >= 1 && j.x_hi >= j.x_lo && j.y_hi >= j.y_lo;
  {
    const int rw = (j.x_hi - j.x_lo) + w, rh = (j.y_hi - j.y_lo) + h;
    ok = ok && j.x_lo >= -a.ref.xorigin && j.y_lo >= -a.ref.yorigin && a.ref.xorigin + j.x_lo + rw <= a.ref.alloc_w &&
         a.ref.yorigin + j.y_lo + rh <= a.ref.alloc_h;
    ok = ok && j.po_x >= -a.org.xorigin && j.po_y >= -a.org.yorigin && a.org.xorigin + j.po_x + w <= a.org.alloc_w &&
         a.org.yorigin + j.po_y + h <= a.org.alloc_h;
  }
  if (!ok) {
    if (tid == 0) {
      r1c_me_result r;
      r.mv.row = 0;
      r.mv.col = 0;
      r.reserved = 1;  // error flag
      r.cost = ~0ull;
      a.results[ji] = r;
    }
    return;
  }

  const int wpr = (w * PB) >> 2;  // words per block row
  {
    const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.po_y * a.org.stride + j.po_x) * PB;
    const long long so = (long long)a.org.stride * PB;
    for (int i = tid; i < wpr * h; i += FS_THREADS) {
      const int r = i / wpr, k = i - r * wpr;
      s_org[i] = me::ldg4_any(po + r * so + 4 * k);
    }
  }
  __syncthreads();

  const int nx = (j.x_hi - j.x_lo) / step + 1, ny = (j.y_hi - j.y_lo) / step + 1;
  const int nc = nx * ny;
  const long long sr = (long long)a.ref.stride * PB;
  Best best;
  best.cost = ~0ull;
  best.idx = 0xffffffffu;
  for (int i = tid; i < nc; i += FS_THREADS) {
    const int iy = i / nx, ix = i - iy * nx;
    const int x = j.x_lo + ix * step, y = j.y_lo + iy * step;
    const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)y * a.ref.stride + x) * PB;
    uint32_t sad = 0;
    for (int r = 0; r < h; r++) {
      const uint8_t *row = pr + r * sr;
      const uint32_t *orow = s_org + r * wpr;
#pragma unroll 4
      for (int k = 0; k < wpr; k++) sad = me::sad_word<PB>(orow[k], me::ldg4_any(row + 4 * k), sad);
    }
    r1c_mv mv;  // src/me.rs:974-977 (i16 arithmetic)
    mv.row = (int16_t)(8 * (int16_t)((int16_t)y - (int16_t)j.po_y));
    mv.col = (int16_t)(8 * (int16_t)((int16_t)x - (int16_t)j.po_x));
    const uint32_t rate = me::mv_rate2(mv, j.pmv[0], j.pmv[1], j.allow_high_precision_mv);
    const unsigned long long cost = 256ull * sad + (unsigned long long)rate * j.lambda;
    if (cost < best.cost) {  // i increases per thread: strict < keeps the first
      best.cost = cost;
      best.idx = (uint32_t)i;
    }
  }
  // lexicographic (cost, idx) min over the CTA
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    const unsigned long long oc = __shfl_xor_sync(0xffffffffu, best.cost, m);
    const uint32_t oi = __shfl_xor_sync(0xffffffffu, best.idx, m);
    if (better(oc, oi, best)) {
      best.cost = oc;
      best.idx = oi;
    }
  }
  if ((tid & 31) == 0) s_best[tid >> 5] = best;
  __syncthreads();
  if (tid == 0) {
    for (int k = 1; k < FS_THREADS / 32; k++)
      if (better(s_best[k].cost, s_best[k].idx, best)) best = s_best[k];
    r1c_me_result out;
    out.reserved = 0;
    if (j.has_init) {
      out = a.results[ji];
      out.reserved = 0;
    } else {
      out.mv.row = 0;
      out.mv.col = 0;
      out.cost = ~0ull;
    }
    if (best.idx != 0xffffffffu && best.cost < out.cost) {  // `cost < *lowest_cost`, src/me.rs:984
      const int iy = (int)best.idx / nx, ix = (int)best.idx - iy * nx;
      const int x = j.x_lo + ix * step, y = j.y_lo + iy * step;
      out.mv.row = (int16_t)(8 * (int16_t)((int16_t)y - (int16_t)j.po_y));
      out.mv.col = (int16_t)(8 * (int16_t)((int16_t)x - (int16_t)j.po_x));
      out.cost = best.cost;
    }
    a.results[ji] = out;
  }
}

int launch_fs(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_fs_job *d_jobs,
              int64_t n, r1c_me_result *d_results) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  FsArgs a;
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.jobs = d_jobs;
  a.results = d_results;
  a.n = n;
  if (org->px_bytes == 1) full_search_kernel<1><<<(unsigned)n, FS_THREADS, 0, ctx->stream>>>(a);
  else full_search_kernel<2><<<(unsigned)n, FS_THREADS, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

int check_planes(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int bit_depth) {
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes, R1C_ERR_ARG, "org/ref pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, org->px_bytes);
  return R1C_OK;
}

}  // namespace

extern "C" void r1c_mv_range(int64_t w_in_b, int64_t h_in_b, int64_t bo_x, int64_t bo_y, int blk_w,
                             int blk_h, int64_t out[4]) {
  // get_mv_range, src/me.rs:64-80 (MI_SIZE = 4)
  const int64_t border_w = 128 + (int64_t)blk_w * 8, border_h = 128 + (int64_t)blk_h * 8;
  out[0] = -bo_x * 32 - border_w;
  out[1] = (w_in_b - bo_x - blk_w / 4) * 32 + border_w;
  out[2] = -bo_y * 32 - border_h;
  out[3] = (h_in_b - bo_y - blk_h / 4) * 32 + border_h;
}

extern "C" uint32_t r1c_mv_rate(r1c_mv a, r1c_mv b, int allow_hp) { return me::mv_rate(a, b, allow_hp); }

extern "C" int r1c_full_search_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                         const r1c_fs_job *d_jobs, int64_t n, int bit_depth,
                                         r1c_me_result *d_results) {
  int rc = check_planes(ctx, org, ref, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 31) && (n == 0 || (d_jobs && d_results)), R1C_ERR_ARG, "bad job/result array");
  return launch_fs(ctx, org, ref, d_jobs, n, d_results);
}

extern "C" int r1c_full_search_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                     const r1c_fs_job *jobs, int64_t n, int bit_depth,
                                     r1c_me_result *results) {
  int rc = check_planes(ctx, org, ref, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 31) && (n == 0 || (jobs && results)), R1C_ERR_ARG, "bad job/result array");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++) {
    const r1c_fs_job &j = jobs[i];
    R1C_REQUIRE(ctx, j.bsize >= 0 && j.bsize < R1C_BLOCK_SIZES_ALL, R1C_ERR_ARG, "fs job %lld: bad block size", (long long)i);
    R1C_REQUIRE(ctx, j.step >= 1 && j.x_hi >= j.x_lo && j.y_hi >= j.y_lo, R1C_ERR_ARG, "fs job %lld: empty window or step < 1", (long long)i);
    const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
    R1C_REQUIRE(ctx, region_in_plane(ref->cfg, j.x_lo, j.y_lo, (int64_t)(j.x_hi - j.x_lo) + w, (int64_t)(j.y_hi - j.y_lo) + h),
                R1C_ERR_BOUNDS, "fs job %lld: search region leaves the allocated plane", (long long)i);
    R1C_REQUIRE(ctx, region_in_plane(org->cfg, j.po_x, j.po_y, w, h), R1C_ERR_BOUNDS, "fs job %lld: org block leaves the allocated plane", (long long)i);
  }
  void *h_jobs, *h_res, *d_jobs, *d_res;
  const size_t jb = (size_t)n * sizeof(r1c_fs_job), rb = (size_t)n * sizeof(r1c_me_result);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, rb, &h_res))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, rb, &d_res))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_res, results, rb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_res, h_res, rb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_fs(ctx, org, ref, (const r1c_fs_job *)d_jobs, n, (r1c_me_result *)d_res))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_res, d_res, rb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(results, h_res, rb);
  return R1C_OK;
}
