// rdo.cu -- RDO distortion kernels:
//   sse_wxh            src/rdo.rs:286-335   SSE per importance block x f64 bias, truncated per block
//   cdef_dist_wxh      src/rdo.rs:219-283   per-8x8 variance / SSIM-boost formula in f64 (IEEE mul/div/sqrt)
//   compute_distortion_bias  src/rdo.rs:477-508   f32 mean of block_importances, (mean / 3) as f64 + 0.65
//   tx-domain SSE      src/encoder.rs:1210-1224
//
// Mapping: one warp per job.  A lane owns one row (<= 8 px) of one importance
// block; the rows of a block sit in `bh` consecutive lanes (bh = 2, 4 or 8) and
// are summed with xor-shuffles inside that segment; the segment's first lane
// does the f64 tail (explicit round-to-nearest intrinsics: the reference never
// fuses a multiply-add) and the u64 block terms are added warp-wide.  Integer
// sums are order-free, the f32 importance sum runs in the reference's (y, x)
// order, so results are bit-identical.
#include "common.cuh"

namespace {

struct RdoArgs {
  PlaneView s1, s2;
  const r1c_rdo_job *jobs;
  long long n;
  uint64_t *out;
  const float *imp;  // nullptr: bias 1.0
  r1c_bias_params bp;
  int xdec, ydec, bit_depth;
};

// Rust `f64 as u64` (saturating, NaN -> 0) == cvt.rzi.u64.f64
__device__ __forceinline__ uint64_t f64_as_u64(double v) { return __double2ull_rz(v); }

__device__ __forceinline__ double distortion_bias(const RdoArgs &a, int bo_x, int bo_y, int wmi, int hmi) {
  if (!a.imp) return 1.0;
  const int x2 = min(bo_x + wmi, a.bp.w_in_b), y2 = min(bo_y + hmi, a.bp.h_in_b);
  float total = 0.f;
  for (int y = bo_y; y < y2; y++)
    for (int x = bo_x; x < x2; x++) {
      const int idx = (y >> 1) * a.bp.w_in_imp_b + (x >> 1);
      total = __fadd_rn(total, idx < a.bp.n_imp ? __ldg(a.imp + idx) : 0.f);
    }
  const float mean = __fdiv_rn(total, (float)(wmi * hmi));
  return __dadd_rn((double)__fdiv_rn(mean, 3.f), 0.65);
}

template <int PB>
__device__ __forceinline__ int px(const PlaneView &v, int x, int y) {
  return PB == 1 ? (int)((const uint8_t *)v.p0)[(long long)y * v.stride + x]
                 : (int)((const uint16_t *)v.p0)[(long long)y * v.stride + x];
}

template <int PB>
__device__ __forceinline__ bool in_plane(const PlaneView &v, int x, int y, int w, int h) {
  return x >= -v.xorigin && y >= -v.yorigin && v.xorigin + x + w <= v.alloc_w && v.yorigin + y + h <= v.alloc_h;
}

// KIND 0: sse_wxh, 1: cdef_dist_wxh
template <int PB, int KIND>
__global__ void __launch_bounds__(256) rdo_dist_kernel(RdoArgs a) {
  const int lane = threadIdx.x & 31;
  const long long job = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= a.n) return;
  const r1c_rdo_job j = a.jobs[job];
  if (!in_plane<PB>(a.s1, j.src_x, j.src_y, j.w, j.h) || !in_plane<PB>(a.s2, j.rec_x, j.rec_y, j.w, j.h)) {
    if (lane == 0) a.out[job] = ~0ull;
    return;
  }
  const int imp_w = KIND ? 8 : min(j.w, 8), imp_h = KIND ? 8 : min(j.h, 8);
  const int bw = KIND ? 8 : imp_w >> a.xdec, bh = KIND ? 8 : imp_h >> a.ydec;  // pixels of this plane
  const int nbx = j.w / bw, nblk = nbx * (j.h / bh);
  const int items = nblk * bh;  // (block, row) pairs; bh is a power of two <= 8
  uint64_t acc = 0;
  for (int it0 = 0; it0 < items; it0 += 32) {  // warp-uniform trip count
    const int it = it0 + lane;
    const bool live = it < items;
    const int blk = it / bh, row = it - blk * bh;
    const int by = blk / nbx, bx = blk - by * nbx;
    uint32_t v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0;
    if (live) {
      const int x1 = j.src_x + bx * bw, y1 = j.src_y + by * bh + row;
      const int x2 = j.rec_x + bx * bw, y2 = j.rec_y + by * bh + row;
      for (int i = 0; i < bw; i++) {
        const int s = px<PB>(a.s1, x1 + i, y1), d = px<PB>(a.s2, x2 + i, y2);
        if (KIND == 0) {
          const int c = s - d;
          v0 += (uint32_t)(c * c);
        } else {  // 64 px x 4095^2 < 2^32: u32 holds every sum of an 8x8 block
          v0 += s;
          v1 += d;
          v2 += s * s;
          v3 += d * d;
          v4 += s * d;
        }
      }
    }
    // sum the rows of each block (bh consecutive lanes); every lane of the warp takes part
    for (int m = 1; m < bh; m <<= 1) {
      v0 += __shfl_xor_sync(0xffffffffu, v0, m);
      if (KIND) {
        v1 += __shfl_xor_sync(0xffffffffu, v1, m);
        v2 += __shfl_xor_sync(0xffffffffu, v2, m);
        v3 += __shfl_xor_sync(0xffffffffu, v3, m);
        v4 += __shfl_xor_sync(0xffffffffu, v4, m);
      }
    }
    if (live && row == 0) {
      uint64_t value;
      if (KIND == 0) {
        value = v0;
      } else {
        const int cs = a.bit_depth - 8;
        const long long ss = v0, sd = v1, s2 = v2, d2 = v3, sdx = v4;
        const double svar = (double)(s2 - ((ss * ss + 32) >> 6));
        const double dvar = (double)(d2 - ((sd * sd + 32) >> 6));
        const double sse = (double)(d2 + s2 - 2 * sdx);
        const double num = __dmul_rn(4033.0 / 16384.0, __dadd_rn(__dadd_rn(svar, dvar), (double)(16384ll << (2 * cs))));
        const double den = __dsqrt_rn(__dadd_rn((double)(16265089ull << (4 * cs)), __dmul_rn(svar, dvar)));
        value = f64_as_u64(__dadd_rn(__dmul_rn(sse, __ddiv_rn(num, den)), 0.5));
      }
      // frame_block_offset of the sub-region (src/tiling/plane_region.rs:264-276)
      const int fbx = (j.src_x + bx * bw) >> (2 - a.xdec), fby = (j.src_y + by * bh) >> (2 - a.ydec);
      const double bias = distortion_bias(a, fbx, fby, imp_w >> 2, imp_h >> 2);
      acc += f64_as_u64(__dmul_rn((double)value, bias));
    }
  }
#pragma unroll
  for (int m = 16; m; m >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
  if (lane == 0) a.out[job] = acc;
}

// one warp per block; coeffs blocks are `area` apart, rcoeffs blocks `coded` apart
__global__ void __launch_bounds__(256) tx_dist_kernel(const int32_t *coeffs, const int32_t *rcoeffs, int area,
                                                      int coded, int bits, long long n, uint64_t *out) {
  const int lane = threadIdx.x & 31;
  const long long b = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= n) return;
  const int32_t *c = coeffs + b * area, *r = rcoeffs + b * coded;
  uint64_t t = 0;
  for (int i = lane; i < coded; i += 32) {
    const uint32_t d = (uint32_t)__ldg(c + i) - (uint32_t)__ldg(r + i);
    t += (uint64_t)(long long)(int32_t)(d * d);  // (c * c) as u64 with the wrapping i32 product
  }
#pragma unroll
  for (int m = 16; m; m >>= 1) t += __shfl_xor_sync(0xffffffffu, t, m);
  if (lane == 0) out[b] = (t + (1ull << (bits - 1))) >> bits;
}

int launch_rdo(r1c_ctx *ctx, int kind, const r1c_plane *s1, const r1c_plane *s2, const r1c_rdo_job *d_jobs, int64_t n,
               int bit_depth, const float *d_imp, const r1c_bias_params *bp, uint64_t *d_out) {
  if (n == 0) return R1C_OK;
  RdoArgs a;
  a.s1 = make_view(s1);
  a.s2 = make_view(s2);
  a.jobs = d_jobs;
  a.n = n;
  a.out = d_out;
  a.imp = d_imp;
  a.bp = bp ? *bp : r1c_bias_params{0, 0, 0, 0};
  a.xdec = (int)s1->cfg.xdec;
  a.ydec = (int)s1->cfg.ydec;
  a.bit_depth = bit_depth;
  const unsigned grid = (unsigned)((n + 7) / 8);
  r1c_prof_begin(ctx, R1C_PROF_DIST);
  if (s1->px_bytes == 1) {
    if (kind) rdo_dist_kernel<1, 1><<<grid, 256, 0, ctx->stream>>>(a);
    else rdo_dist_kernel<1, 0><<<grid, 256, 0, ctx->stream>>>(a);
  } else {
    if (kind) rdo_dist_kernel<2, 1><<<grid, 256, 0, ctx->stream>>>(a);
    else rdo_dist_kernel<2, 0><<<grid, 256, 0, ctx->stream>>>(a);
  }
  r1c_prof_end(ctx, R1C_PROF_DIST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

int check_rdo(r1c_ctx *ctx, int kind, const r1c_plane *s1, const r1c_plane *s2, int64_t n, int bit_depth,
              const float *imp, const r1c_bias_params *bp) {
  if (!ctx || !s1 || !s2) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, kind == 0 || kind == 1, R1C_ERR_ARG, "kind must be 0 (sse_wxh) or 1 (cdef_dist_wxh)");
  R1C_REQUIRE(ctx, s1->px_bytes == s2->px_bytes, R1C_ERR_ARG, "src1/src2 pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, s1->px_bytes), R1C_ERR_ARG, "bit_depth %d does not match %d-byte pixels", bit_depth, s1->px_bytes);
  R1C_REQUIRE(ctx, s1->cfg.xdec >= 0 && s1->cfg.xdec <= 1 && s1->cfg.ydec >= 0 && s1->cfg.ydec <= 1, R1C_ERR_ARG, "decimation must be 0 or 1");
  // cdef_dist is only defined on non-subsampled planes (debug_asserts of src/rdo.rs:222-225)
  R1C_REQUIRE(ctx, kind == 0 || (s1->cfg.xdec == 0 && s1->cfg.ydec == 0 && s2->cfg.xdec == 0 && s2->cfg.ydec == 0), R1C_ERR_ARG,
              "cdef_dist_wxh needs non-subsampled planes");
  R1C_REQUIRE(ctx, n >= 0, R1C_ERR_ARG, "negative job count");
  R1C_REQUIRE(ctx, !imp || (bp && bp->n_imp > 0 && bp->w_in_imp_b > 0), R1C_ERR_ARG, "block_importances without its geometry");
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_rdo_dist_batch_dev(r1c_ctx *ctx, int kind, const r1c_plane *src1, const r1c_plane *src2,
                                      const r1c_rdo_job *d_jobs, int64_t n, int bit_depth,
                                      const float *d_block_importances, const r1c_bias_params *bp, uint64_t *d_out) {
  int rc = check_rdo(ctx, kind, src1, src2, n, bit_depth, d_block_importances, bp);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (d_jobs && d_out), R1C_ERR_ARG, "null job/result array");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_rdo(ctx, kind, src1, src2, d_jobs, n, bit_depth, d_block_importances, bp, d_out);
}

static int rdo_batch_host(r1c_ctx *ctx, int kind, const r1c_plane *s1, const r1c_plane *s2, const r1c_rdo_job *jobs,
                          int64_t n, int bit_depth, const float *imp, const r1c_bias_params *bp, uint64_t *out) {
  int rc = check_rdo(ctx, kind, s1, s2, n, bit_depth, imp, bp);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (jobs && out), R1C_ERR_ARG, "null job/result array");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++) {
    const r1c_rdo_job &j = jobs[i];
    // the asserts of sse_wxh / cdef_dist_wxh (src/rdo.rs:260-261,290-291) + region bounds
    const int mask = kind ? 7 : 3;
    R1C_REQUIRE(ctx, j.w > 0 && j.h > 0 && j.w <= 128 && j.h <= 128 && !(j.w & mask) && !(j.h & mask), R1C_ERR_ARG,
                "job %lld: %dx%d is not a multiple of %d", (long long)i, j.w, j.h, mask + 1);
    R1C_REQUIRE(ctx, region_in_plane(s1->cfg, j.src_x, j.src_y, j.w, j.h) && region_in_plane(s2->cfg, j.rec_x, j.rec_y, j.w, j.h),
                R1C_ERR_BOUNDS, "job %lld: block leaves the allocated plane", (long long)i);
    R1C_REQUIRE(ctx, j.src_x >= 0 && j.src_y >= 0, R1C_ERR_ARG, "job %lld: negative region origin has no frame block offset", (long long)i);
  }
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *h_jobs, *d_jobs, *d_out, *h_out, *d_imp = nullptr, *h_imp;
  const size_t jb = (size_t)n * sizeof(r1c_rdo_job), ob = (size_t)n * sizeof(uint64_t);
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, ob, &h_out))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  memcpy(h_jobs, jobs, jb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  if (imp) {
    const size_t ib = (size_t)bp->n_imp * sizeof(float);
    if ((rc = r1c_stage_host(ctx, 2, ib, &h_imp))) return rc;
    if ((rc = r1c_scratch_dev(ctx, 2, ib, &d_imp))) return rc;
    memcpy(h_imp, imp, ib);
    R1C_CUDA(ctx, cudaMemcpyAsync(d_imp, h_imp, ib, cudaMemcpyHostToDevice, ctx->stream));
  }
  if ((rc = launch_rdo(ctx, kind, s1, s2, (const r1c_rdo_job *)d_jobs, n, bit_depth, (const float *)d_imp, bp, (uint64_t *)d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, h_out, ob);
  return R1C_OK;
}

extern "C" int r1c_sse_wxh_batch(r1c_ctx *ctx, const r1c_plane *src1, const r1c_plane *src2, const r1c_rdo_job *jobs,
                                 int64_t n, int bit_depth, const float *block_importances, const r1c_bias_params *bp,
                                 uint64_t *out) {
  return rdo_batch_host(ctx, 0, src1, src2, jobs, n, bit_depth, block_importances, bp, out);
}
extern "C" int r1c_cdef_dist_batch(r1c_ctx *ctx, const r1c_plane *src1, const r1c_plane *src2, const r1c_rdo_job *jobs,
                                   int64_t n, int bit_depth, const float *block_importances, const r1c_bias_params *bp,
                                   uint64_t *out) {
  return rdo_batch_host(ctx, 1, src1, src2, jobs, n, bit_depth, block_importances, bp, out);
}

extern "C" int r1c_tx_dist_batch_dev(r1c_ctx *ctx, const int32_t *d_coeffs, const int32_t *d_rcoeffs, int tx_size,
                                     int64_t n, uint64_t *d_out) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_coeffs && d_rcoeffs && d_out)), R1C_ERR_ARG, "null coefficient/result array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size), area = w * h;
  const int coded = (w < 32 ? w : 32) * (h < 32 ? h : 32);      // av1_get_coded_tx_size, src/encoder.rs:1143
  const int bits = 2 * (3 - ((area > 256) + (area > 1024)));    // get_log_tx_scale, src/quantize.rs:34-39
  r1c_prof_begin(ctx, R1C_PROF_DIST);
  tx_dist_kernel<<<(unsigned)((n + 7) / 8), 256, 0, ctx->stream>>>(d_coeffs, d_rcoeffs, area, coded, bits, n, d_out);
  r1c_prof_end(ctx, R1C_PROF_DIST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_tx_dist_batch(r1c_ctx *ctx, const int32_t *coeffs, const int32_t *rcoeffs, int tx_size, int64_t n,
                                 uint64_t *out) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, tx_size >= 0 && tx_size < R1C_TX_SIZES_ALL, R1C_ERR_ARG, "bad tx_size %d", tx_size);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (coeffs && rcoeffs && out)), R1C_ERR_ARG, "null coefficient/result array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const int w = r1c_tx_width(tx_size), h = r1c_tx_height(tx_size);
  const size_t cb = (size_t)n * w * h * 4, rb = (size_t)n * (w < 32 ? w : 32) * (h < 32 ? h : 32) * 4, ob = (size_t)n * 8;
  void *d_c, *d_r, *d_o, *h_o;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 0, cb, &d_c))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, rb, &d_r))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, ob, &d_o))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, ob, &h_o))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_c, coeffs, cb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_r, rcoeffs, rb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_tx_dist_batch_dev(ctx, (const int32_t *)d_c, (const int32_t *)d_r, tx_size, n, (uint64_t *)d_o))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_o, d_o, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, h_o, ob);
  return R1C_OK;
}
