// dist.cu -- batched SAD / SATD (get_sad / get_satd, reference src/dist.rs:48-188;
// semantics of get_sad_ref :25-46 and get_satd_ref :197-328).
//
// Layout: one warp per job.  Pixels are fetched as aligned 32-bit words and
// re-aligned with a funnel shift, so a block may sit at any x in either plane.
//   SAD  u8 : VABSDIFF4.U8.ACC (4 abs-diffs + accumulate in one instruction)
//   SAD  u16: two 16-bit lanes per word, IABS + IADD
//   SATD    : lane = one row of one 4x4 / 8x8 tile; horizontal Hadamard in
//             registers, vertical Hadamard with __shfl_xor butterflies.  The
//             Hadamard is exact integer arithmetic and the result is a sum of
//             |coef| over the tile, so the pass order / output permutation of
//             the reference (vertical first, src/dist.rs:259-272) cannot change
//             the value.
// HBM-bound: 2*px_bytes bytes per pixel + 4 B per job.
#include "dist_dev.cuh"

namespace {
using namespace distdev;

struct DistArgs {
  PlaneView org, ref;
  const r1c_dist_job *jobs;
  uint32_t *out;
  long long n;
};

__device__ __forceinline__ bool job_in_bounds(const PlaneView &v, int x, int y, int w, int h) {
  return x >= -v.xorigin && y >= -v.yorigin && v.xorigin + x + w <= v.alloc_w &&
         v.yorigin + y + h <= v.alloc_h;
}

template <int PB>
__global__ void __launch_bounds__(256) sad_kernel(DistArgs a) {
  const int lane = threadIdx.x & 31;
  long long job = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= a.n) return;
  const r1c_dist_job j = a.jobs[job];
  const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
  if (!job_in_bounds(a.org, j.org_x, j.org_y, w, h) || !job_in_bounds(a.ref, j.ref_x, j.ref_y, w, h)) {
    if (lane == 0) a.out[job] = 0xFFFFFFFFu;
    return;
  }
  const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.org_y * a.org.stride + j.org_x) * PB;
  const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)j.ref_y * a.ref.stride + j.ref_x) * PB;
  const int so = a.org.stride * PB, sr = a.ref.stride * PB;
  const uint32_t acc = sad_warp<PB>(po, so, pr, sr, w, h, lane);
  if (lane == 0) a.out[job] = acc;
}

template <int PB>
__global__ void __launch_bounds__(256) satd_kernel(DistArgs a) {
  const int lane = threadIdx.x & 31;
  long long job = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= a.n) return;
  const r1c_dist_job j = a.jobs[job];
  const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
  if (!job_in_bounds(a.org, j.org_x, j.org_y, w, h) || !job_in_bounds(a.ref, j.ref_x, j.ref_y, w, h)) {
    if (lane == 0) a.out[job] = 0xFFFFFFFFu;
    return;
  }
  const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.org_y * a.org.stride + j.org_x) * PB;
  const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)j.ref_y * a.ref.stride + j.ref_x) * PB;
  const int so = a.org.stride * PB, sr = a.ref.stride * PB;
  const uint32_t v = satd_block_warp<PB>(po, so, pr, sr, w, h, lane);
  if (lane == 0) a.out[job] = v;
}

int launch_dist(r1c_ctx *ctx, int use_satd, const r1c_plane *org, const r1c_plane *ref,
                const r1c_dist_job *d_jobs, int64_t n, uint32_t *d_out) {
  if (n == 0) return R1C_OK;
  DistArgs a;
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.jobs = d_jobs;
  a.out = d_out;
  a.n = n;
  const int wpb = 8;
  unsigned grid = (unsigned)((n + wpb - 1) / wpb);
  r1c_prof_begin(ctx, R1C_PROF_DIST);
  if (org->px_bytes == 1) {
    if (use_satd) satd_kernel<1><<<grid, wpb * 32, 0, ctx->stream>>>(a);
    else sad_kernel<1><<<grid, wpb * 32, 0, ctx->stream>>>(a);
  } else {
    if (use_satd) satd_kernel<2><<<grid, wpb * 32, 0, ctx->stream>>>(a);
    else sad_kernel<2><<<grid, wpb * 32, 0, ctx->stream>>>(a);
  }
  r1c_prof_end(ctx, R1C_PROF_DIST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

int check_common(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, int bit_depth) {
  if (!ctx || !org || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes, R1C_ERR_ARG, "org/ref pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG, "bit_depth %d does not match %d-byte pixels", bit_depth, org->px_bytes);
  return R1C_OK;
}

int dist_batch_host(r1c_ctx *ctx, int use_satd, const r1c_plane *org, const r1c_plane *ref,
                    const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out) {
  int rc = check_common(ctx, org, ref, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (jobs && out)), R1C_ERR_ARG, "null job/result array");
  if (n == 0) return R1C_OK;
  // the asserts of get_sad / get_satd (src/dist.rs:54-69) + PlaneRegion bounds
  for (int64_t i = 0; i < n; i++) {
    const r1c_dist_job &j = jobs[i];
    R1C_REQUIRE(ctx, j.bsize >= 0 && j.bsize < R1C_BLOCK_SIZES_ALL, R1C_ERR_ARG, "job %lld: bad block size %d", (long long)i, j.bsize);
    int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
    R1C_REQUIRE(ctx, region_in_plane(org->cfg, j.org_x, j.org_y, w, h) && region_in_plane(ref->cfg, j.ref_x, j.ref_y, w, h),
                R1C_ERR_BOUNDS, "job %lld: block leaves the allocated plane", (long long)i);
  }
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *h_jobs, *d_jobs, *d_out, *h_out;
  size_t jb = (size_t)n * sizeof(r1c_dist_job), ob = (size_t)n * sizeof(uint32_t);
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, ob, &h_out))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  memcpy(h_jobs, jobs, jb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_dist(ctx, use_satd, org, ref, (const r1c_dist_job *)d_jobs, n, (uint32_t *)d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, h_out, ob);
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_sad_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                             const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out) {
  return dist_batch_host(ctx, 0, org, ref, jobs, n, bit_depth, out);
}
extern "C" int r1c_satd_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                              const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out) {
  return dist_batch_host(ctx, 1, org, ref, jobs, n, bit_depth, out);
}
extern "C" int r1c_dist_batch_dev(r1c_ctx *ctx, int use_satd, const r1c_plane *org,
                                  const r1c_plane *ref, const r1c_dist_job *d_jobs, int64_t n,
                                  int bit_depth, uint32_t *d_out) {
  int rc = check_common(ctx, org, ref, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_out)), R1C_ERR_ARG, "null job/result array");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_dist(ctx, use_satd, org, ref, d_jobs, n, d_out);
}
extern "C" int r1c_sad(r1c_ctx *ctx, const r1c_plane *org, int ox, int oy, const r1c_plane *ref,
                       int rx, int ry, int bsize, int bit_depth, uint32_t *out) {
  r1c_dist_job j = {ox, oy, rx, ry, bsize};
  return dist_batch_host(ctx, 0, org, ref, &j, 1, bit_depth, out);
}
extern "C" int r1c_satd(r1c_ctx *ctx, const r1c_plane *org, int ox, int oy, const r1c_plane *ref,
                        int rx, int ry, int bsize, int bit_depth, uint32_t *out) {
  r1c_dist_job j = {ox, oy, rx, ry, bsize};
  return dist_batch_host(ctx, 1, org, ref, &j, 1, bit_depth, out);
}
