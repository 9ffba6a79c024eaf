// dist.cu -- batched SAD / SATD (get_sad / get_satd, reference src/dist.rs:48-188;
// semantics of get_sad_ref :25-46 and get_satd_ref :197-328).
//
// Layout: one warp per job.  Pixels are fetched as aligned 32-bit words and
// re-aligned with a funnel shift, so a block may sit at any x in either plane.
//   SAD  u8 : VABSDIFF4.U8.ACC (4 abs-diffs + accumulate in one instruction)
//   SAD  u16: two 16-bit lanes per word, max - min per half (VIMNMX.U16x2) + IDP.2A
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

// ---- lane-per-tile kernels ------------------------------------------------------------
// A lane owns whole S x S tiles (S = 4 when the block has a side of 4, else 8): it loads
// the S rows of both planes with word loads, and for SATD keeps the S*S differences in
// registers, where the 2-D Hadamard needs no shuffles (6 add/sub per pixel).  G lanes
// share a block and stride over its tiles; G is picked from the length of the job list
// (long lists = small blocks = one lane per block, so a warp works on 32 blocks and the
// per-block set-up is paid once per warp instead of once per 64 pixels).

template <int PB, int S, bool VEC>
__device__ __forceinline__ uint32_t tile_sad(const uint8_t *po, int so, const uint8_t *pr, int sr) {
  constexpr int NW = S * PB / 4;
  // every row of both tiles is requested before the first is used (with a load -> use -> load chain per row the
  // kernel ran at 48 registers and one row of memory latency at a time: slower than the SATD of the same tiles)
  uint32_t x[S][NW], y[S][NW];
#pragma unroll
  for (int r = 0; r < S; r++) {
    ld_row<NW, VEC>(po + r * so, x[r]);
    ld_row<NW, VEC>(pr + r * sr, y[r]);
  }
  uint32_t acc = 0;
#pragma unroll
  for (int r = 0; r < S; r++) {
#pragma unroll
    for (int i = 0; i < NW; i++) acc = PB == 1 ? sad4_acc(x[r][i], y[r][i], acc) : sad2_acc(x[r][i], y[r][i], acc);
  }
  return acc;
}

template <int PB, int S, bool VEC>
__device__ __forceinline__ uint32_t tile_satd(const uint8_t *po, int so, const uint8_t *pr, int sr) {
  int d[S][S];
#pragma unroll
  for (int r = 0; r < S; r++) {
    load_diff_row<PB, S, VEC>(po + r * so, pr + r * sr, d[r]);
    hadamard_regs<S>(d[r]);  // horizontal
  }
#pragma unroll
  for (int m = 1; m < S; m <<= 1)  // vertical
#pragma unroll
    for (int r = 0; r < S; r++)
      if ((r & m) == 0) {
#pragma unroll
        for (int c = 0; c < S; c++) {
          const int u = d[r][c], v = d[r | m][c];
          d[r][c] = u + v;
          d[r | m][c] = u - v;
        }
      }
  uint32_t acc = 0;
#pragma unroll
  for (int r = 0; r < S; r++)
#pragma unroll
    for (int c = 0; c < S; c++) acc += (uint32_t)abs(d[r][c]);
  return acc;
}

template <int PB, int G, int SATD, bool VEC>
__global__ void __launch_bounds__(128) dist_kernel(DistArgs a) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long job = gid / G;
  const int sub = (int)(gid % G);
  const unsigned mask = G == 32 ? 0xffffffffu : (G == 1 ? 0u : (((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1))));
  if (job >= a.n) return;
  const r1c_dist_job j = a.jobs[job];
  const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
  if (!job_in_bounds(a.org, j.org_x, j.org_y, w, h) || !job_in_bounds(a.ref, j.ref_x, j.ref_y, w, h)) {
    if (sub == 0) a.out[job] = 0xFFFFFFFFu;
    return;
  }
  const uint8_t *po = (const uint8_t *)a.org.p0 + ((long long)j.org_y * a.org.stride + j.org_x) * PB;
  const uint8_t *pr = (const uint8_t *)a.ref.p0 + ((long long)j.ref_y * a.ref.stride + j.ref_x) * PB;
  const int so = a.org.stride * PB, sr = a.ref.stride * PB;
  const bool small = w == 4 || h == 4;  // get_satd_ref's tile rule (src/dist.rs:287); SAD uses the same tiling
  const int ls = small ? 2 : 3;
  const int tx = w >> ls, ntiles = tx * (h >> ls);
  const int lg_tx = __ffs(tx) - 1;  // BlockSize widths are powers of two
  unsigned long long acc = 0;  // 256 tiles x 8*8*64*4095 exceeds u32
  for (int t = sub; t < ntiles; t += G) {
    const int ty = t >> lg_tx, txi = t - (ty << lg_tx);
    const uint8_t *o = po + (long long)(ty << ls) * so + (txi << ls) * PB;
    const uint8_t *r = pr + (long long)(ty << ls) * sr + (txi << ls) * PB;
    if (small) acc += SATD ? tile_satd<PB, 4, VEC>(o, so, r, sr) : tile_sad<PB, 4, VEC>(o, so, r, sr);
    else acc += SATD ? tile_satd<PB, 8, VEC>(o, so, r, sr) : tile_sad<PB, 8, VEC>(o, so, r, sr);
  }
  if (G > 1) {
#pragma unroll
    for (int m = G >> 1; m; m >>= 1) acc += __shfl_xor_sync(mask, acc, m);
  }
  if (sub == 0) a.out[job] = SATD ? (uint32_t)((acc + ((1ull << ls) >> 1)) >> ls) : (uint32_t)acc;
}

template <int PB, int G>
void launch_dist_g(r1c_ctx *ctx, int use_satd, bool vec, const DistArgs &a) {
  const unsigned grid = (unsigned)((a.n * G + 127) / 128);
  if (vec) {
    if (use_satd) dist_kernel<PB, G, 1, true><<<grid, 128, 0, ctx->stream>>>(a);
    else dist_kernel<PB, G, 0, true><<<grid, 128, 0, ctx->stream>>>(a);
  } else {
    if (use_satd) dist_kernel<PB, G, 1, false><<<grid, 128, 0, ctx->stream>>>(a);
    else dist_kernel<PB, G, 0, false><<<grid, 128, 0, ctx->stream>>>(a);
  }
}

template <int PB>
void launch_dist_pb(r1c_ctx *ctx, int use_satd, bool vec, const DistArgs &a) {
  // lanes per block from the list length: enough lanes to fill the GPU (~150 K threads), and a long
  // list means small blocks (a frame tiled by 8x8 has >= 30 K of them), which one lane handles whole
  const long long want = (long long)ctx->sm_count * 1024 / (a.n > 0 ? a.n : 1);
  if (want < 2) launch_dist_g<PB, 1>(ctx, use_satd, vec, a);
  else if (want < 4) launch_dist_g<PB, 2>(ctx, use_satd, vec, a);
  else if (want < 8) launch_dist_g<PB, 4>(ctx, use_satd, vec, a);
  else if (want < 16) launch_dist_g<PB, 8>(ctx, use_satd, vec, a);
  else if (want < 32) launch_dist_g<PB, 16>(ctx, use_satd, vec, a);
  else launch_dist_g<PB, 32>(ctx, use_satd, vec, a);
}

// vector loads are legal when both allocations start and end on 16-byte boundaries
static bool planes_vec_ok(const r1c_plane *p) {
  const size_t bytes = (size_t)p->cfg.stride * p->cfg.alloc_height * p->px_bytes;
  return (((uintptr_t)p->d_data) & 15) == 0 && (bytes & 15) == 0;
}

// compute_lookahead_intra_costs, src/api/internal.rs:680-765: SATD of every 8x8 importance block of the
// input luma plane against its DC_PRED prediction.  The reference passes the block's own rectangle as
// `tile_rect` (:730-735), so predict_intra sees x = y = 0, picks PredictionVariant::NONE (src/predict.rs:176-184,
// 206-212) and the prediction is pred_dc_128 for EVERY block: the constant 128 << (bit_depth - 8) (:623-633).
// One lane per block (the lane-per-tile SATD above with a constant second operand).
template <int PB, bool VEC>
__global__ void __launch_bounds__(128) intra_cost_kernel(PlaneView v, int w_imp, int h_imp, int base, uint32_t *out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= w_imp * h_imp) return;
  const int by = i / w_imp, bx = i - by * w_imp;
  const uint8_t *p = (const uint8_t *)v.p0 + ((long long)(by * 8) * v.stride + bx * 8) * PB;
  const int sb = v.stride * PB;
  int d[8][8];
#pragma unroll
  for (int r = 0; r < 8; r++) {
    constexpr int NW = 8 * PB / 4;
    uint32_t x[NW];
    ld_row<NW, VEC>(p + r * sb, x);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t w = x[k * PB / 4];
      const int px = PB == 1 ? (int)((w >> (8 * (k & 3))) & 0xff) : (int)((w >> (16 * (k & 1))) & 0xffff);
      d[r][k] = px - base;
    }
    hadamard_regs<8>(d[r]);
  }
#pragma unroll
  for (int m = 1; m < 8; m <<= 1)
#pragma unroll
    for (int r = 0; r < 8; r++)
      if ((r & m) == 0) {
#pragma unroll
        for (int c = 0; c < 8; c++) {
          const int u = d[r][c], w2 = d[r | m][c];
          d[r][c] = u + w2;
          d[r | m][c] = u - w2;
        }
      }
  uint32_t acc = 0;
#pragma unroll
  for (int r = 0; r < 8; r++)
#pragma unroll
    for (int c = 0; c < 8; c++) acc += (uint32_t)abs(d[r][c]);
  out[i] = (acc + 4) >> 3;
}

// ---- SATD as the fork's GENERATED kernels compute it: i16 lanes for every pixel type ------------------------
// (build/kernel/gen/dist.rs:171-350, normalisation :553-556; `let calc_prim = PrimType::I16` at :189-194 and :252-257).
// Adds and subtractions mod 2^16 form a ring, so a lane of the generated code holds the exact Hadamard term reduced
// mod 2^16; only the merged last stage -- a.abs().max(b.abs()) * 2 on i16, pairs (k, k + 4) of the two half transforms
// of the second dimension -- the sign-extending cast to u32 and the wrapping u32 sum differ from get_satd_ref.  A
// compatibility switch for comparisons with a reference binary, not a fast path: one warp per job, a lane per tile.
__device__ __forceinline__ int w16(int v) { return (int)(short)v; }
__device__ __forceinline__ uint32_t merged16(int a, int b) {
  const int x = w16(abs(w16(a))), y = w16(abs(w16(b)));
  return (uint32_t)w16(2 * max(x, y));
}

template <typename Px>
__global__ void __launch_bounds__(128) satd_i16_kernel(const __grid_constant__ DistArgs a) {
  const long long job = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= a.n) return;
  const int lane = threadIdx.x & 31;
  const r1c_dist_job j = a.jobs[job];
  const int w = blk_w_of(j.bsize), h = blk_h_of(j.bsize);
  const int S = min(min(w, h), 8), tw = w / S, tiles = tw * (h / S);
  const Px *po = (const Px *)a.org.p0 + (long long)j.org_y * a.org.stride + j.org_x;
  const Px *pr = (const Px *)a.ref.p0 + (long long)j.ref_y * a.ref.stride + j.ref_x;
  uint32_t sum = 0;
  for (int t = lane; t < tiles; t += 32) {
    const int ty = (t / tw) * S, tx = (t % tw) * S;
    const Px *o = po + (long long)ty * a.org.stride + tx, *r = pr + (long long)ty * a.ref.stride + tx;
    if (S == 4) {
      int d[4][4];
#pragma unroll
      for (int y = 0; y < 4; y++)
#pragma unroll
        for (int x = 0; x < 4; x++) d[y][x] = (int)o[(long long)y * a.org.stride + x] - (int)r[(long long)y * a.ref.stride + x];
#pragma unroll
      for (int x = 0; x < 4; x++) {  // vertical 4-point Hadamard, exact
        const int a0 = d[0][x] + d[1][x], a1 = d[0][x] - d[1][x], a2 = d[2][x] + d[3][x], a3 = d[2][x] - d[3][x];
        d[0][x] = a0 + a2, d[1][x] = a1 + a3, d[2][x] = a0 - a2, d[3][x] = a1 - a3;
      }
#pragma unroll
      for (int k = 0; k < 4; k++) sum += merged16(d[k][0] + d[k][1], d[k][2] + d[k][3]) + merged16(d[k][0] - d[k][1], d[k][2] - d[k][3]);
    } else {
      int d[8][8];
#pragma unroll
      for (int y = 0; y < 8; y++)
#pragma unroll
        for (int x = 0; x < 8; x++) d[y][x] = (int)o[(long long)y * a.org.stride + x] - (int)r[(long long)y * a.ref.stride + x];
#pragma unroll
      for (int x = 0; x < 8; x++) {  // vertical 8-point Hadamard, exact (any order: only the sum over a row's terms matters)
#pragma unroll
        for (int st = 1; st < 8; st <<= 1)
#pragma unroll
          for (int y = 0; y < 8; y++)
            if (!(y & st)) {
              const int u = d[y][x], v = d[y | st][x];
              d[y][x] = u + v, d[y | st][x] = u - v;
            }
      }
#pragma unroll
      for (int y = 0; y < 8; y++) {
#pragma unroll
        for (int hh = 0; hh < 8; hh += 4) {  // 4-point transforms of columns 0-3 and 4-7
          const int a0 = d[y][hh] + d[y][hh + 1], a1 = d[y][hh] - d[y][hh + 1], a2 = d[y][hh + 2] + d[y][hh + 3], a3 = d[y][hh + 2] - d[y][hh + 3];
          d[y][hh] = a0 + a2, d[y][hh + 1] = a1 + a3, d[y][hh + 2] = a0 - a2, d[y][hh + 3] = a1 - a3;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) sum += merged16(d[y][k], d[y][4 + k]);
      }
    }
  }
#pragma unroll
  for (int m = 16; m; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
  const int ln = S == 4 ? 2 : 3;
  if (lane == 0) a.out[job] = (sum + ((1u << ln) >> 1)) >> ln;
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
  r1c_prof_begin(ctx, R1C_PROF_DIST);
  const bool vec = planes_vec_ok(org) && planes_vec_ok(ref);
  if (use_satd == R1C_DIST_SATD_I16) {
    const unsigned grid = (unsigned)((n + 3) / 4);
    if (org->px_bytes == 1) satd_i16_kernel<uint8_t><<<grid, 128, 0, ctx->stream>>>(a);
    else satd_i16_kernel<uint16_t><<<grid, 128, 0, ctx->stream>>>(a);
  } else if (org->px_bytes == 1) launch_dist_pb<1>(ctx, use_satd, vec, a);
  else launch_dist_pb<2>(ctx, use_satd, vec, a);
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
  const void *h_jobs;
  void *d_jobs, *d_out, *h_out;
  int staged;
  size_t jb = (size_t)n * sizeof(r1c_dist_job), ob = (size_t)n * sizeof(uint32_t);
  if ((rc = r1c_host_src(ctx, 0, jobs, jb, &h_jobs))) return rc;  // page-locked buffers: DMA source / target themselves
  if ((rc = r1c_host_dst(ctx, 1, out, ob, &h_out, &staged))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_dist(ctx, use_satd, org, ref, (const r1c_dist_job *)d_jobs, n, (uint32_t *)d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (staged) memcpy(out, h_out, ob);
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
extern "C" int r1c_satd_i16_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                  const r1c_dist_job *jobs, int64_t n, int bit_depth, uint32_t *out) {
  return dist_batch_host(ctx, R1C_DIST_SATD_I16, org, ref, jobs, n, bit_depth, out);
}
extern "C" int r1c_dist_batch_dev(r1c_ctx *ctx, int use_satd, const r1c_plane *org,
                                  const r1c_plane *ref, const r1c_dist_job *d_jobs, int64_t n,
                                  int bit_depth, uint32_t *d_out) {
  int rc = check_common(ctx, org, ref, bit_depth);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_out)), R1C_ERR_ARG, "null job/result array");
  R1C_REQUIRE(ctx, use_satd >= R1C_DIST_SAD && use_satd <= R1C_DIST_SATD_I16, R1C_ERR_ARG, "bad distortion kind %d", use_satd);
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

extern "C" int r1c_lookahead_intra_costs_dev(r1c_ctx *ctx, const r1c_plane *input, int w_in_imp_b, int h_in_imp_b,
                                             int bit_depth, uint32_t *d_out) {
  if (!ctx || !input) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, input->px_bytes), R1C_ERR_ARG, "bit_depth %d does not match %d-byte pixels", bit_depth, input->px_bytes);
  R1C_REQUIRE(ctx, w_in_imp_b >= 0 && h_in_imp_b >= 0 && (long long)w_in_imp_b * h_in_imp_b < (1ll << 30) && d_out, R1C_ERR_ARG, "bad importance-block grid");
  R1C_REQUIRE(ctx, region_in_plane(input->cfg, 0, 0, (int64_t)w_in_imp_b * 8, (int64_t)h_in_imp_b * 8), R1C_ERR_BOUNDS,
              "the importance-block grid leaves the allocated plane");
  const int n = w_in_imp_b * h_in_imp_b;
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const PlaneView v = make_view(input);
  const int base = 128 << (bit_depth - 8);
  const bool vec = planes_vec_ok(input);
  const unsigned grid = (unsigned)((n + 127) / 128);
  r1c_prof_begin(ctx, R1C_PROF_DIST);
  if (input->px_bytes == 1) {
    if (vec) intra_cost_kernel<1, true><<<grid, 128, 0, ctx->stream>>>(v, w_in_imp_b, h_in_imp_b, base, d_out);
    else intra_cost_kernel<1, false><<<grid, 128, 0, ctx->stream>>>(v, w_in_imp_b, h_in_imp_b, base, d_out);
  } else {
    if (vec) intra_cost_kernel<2, true><<<grid, 128, 0, ctx->stream>>>(v, w_in_imp_b, h_in_imp_b, base, d_out);
    else intra_cost_kernel<2, false><<<grid, 128, 0, ctx->stream>>>(v, w_in_imp_b, h_in_imp_b, base, d_out);
  }
  r1c_prof_end(ctx, R1C_PROF_DIST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_lookahead_intra_costs(r1c_ctx *ctx, const r1c_plane *input, int w_in_imp_b, int h_in_imp_b, int bit_depth,
                                         uint32_t *out) {
  if (!ctx || !input) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, w_in_imp_b >= 0 && h_in_imp_b >= 0 && (long long)w_in_imp_b * h_in_imp_b < (1ll << 30) && out, R1C_ERR_ARG, "bad importance-block grid");
  const size_t ob = (size_t)w_in_imp_b * h_in_imp_b * sizeof(uint32_t);
  if (ob == 0) return R1C_OK;
  void *d_out, *h_out;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 1, ob, &d_out))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, ob, &h_out))) return rc;
  if ((rc = r1c_lookahead_intra_costs_dev(ctx, input, w_in_imp_b, h_in_imp_b, bit_depth, (uint32_t *)d_out))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, h_out, ob);
  return R1C_OK;
}
