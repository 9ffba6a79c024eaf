// importance.cu -- block-importance propagation of the lookahead
// (reference ContextInner::compute_block_importances, src/api/internal.rs:823-1075).
//
// One call = one (frame, reference) pair of the reference's loop :897-1047: the inter cost of every
// 8x8 importance block (get_satd against the reference plane at the block's lookahead MV), the
// propagated amount, and its bilinear split over the four importance blocks of the reference that
// the displaced block covers.  The reference adds those shares into the reference frame's f32 array
// in raster order of the source blocks; f32 addition is not associative, so the order is kept:
//   1. imp_jobs_kernel      MV -> SATD job (r1c_dist_batch_dev computes the inter costs)
//   2. imp_shares_kernel    4 (target cell, share) pairs per source block, key = target cell
//   3. cub radix sort by key (stable: shares of one target stay in source order)
//   4. imp_accumulate_kernel one thread per target segment adds its shares in that order (__fadd_rn)
// All f32 operations are explicit round-to-nearest intrinsics (no contraction into FMAs).
// r1c_importance_finalize is :1054-1071: log2(1 + importance / intra_cost); log2 is evaluated in f64
// and rounded once.  The reference's value comes from the libm it is linked with (log2f, not correctly
// rounded, platform-dependent): agreement is within 1 ulp of f32 -- the only tolerance of the library.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int IMP_B = 8;              // IMPORTANCE_BLOCK_SIZE, src/api/internal.rs:24
constexpr long long B_MV = IMP_B * 8; // BLOCK_SIZE_IN_MV_UNITS
constexpr uint32_t NO_TARGET = 0x7fffffffu;

struct ImpArgs {
  PlaneView ref;
  const r1c_mv *mvs;
  int mv_cols;
  const uint32_t *intra, *inter;
  const float *imp;
  int n_refs, w, h;
  r1c_dist_job *jobs;
  uint32_t *keys;
  float *vals;
};

__device__ __forceinline__ long long tdiv(long long a, long long b) { return a / b; }  // Rust `/`: toward zero

__global__ void imp_jobs_kernel(ImpArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.w * a.h) return;
  const int y = i / a.w, x = i - y * a.w;
  const r1c_mv mv = a.mvs[(long long)(2 * y) * a.mv_cols + 2 * x];
  const long long rx = tdiv((long long)x * B_MV + mv.col, 8), ry = tdiv((long long)y * B_MV + mv.row, 8);
  r1c_dist_job j;
  j.org_x = x * IMP_B;
  j.org_y = y * IMP_B;
  // clamped into the allocation for memory safety only: the host entry point rejects such MVs beforehand
  j.ref_x = (int)min(max(rx, (long long)-a.ref.xorigin), (long long)(a.ref.alloc_w - a.ref.xorigin - IMP_B));
  j.ref_y = (int)min(max(ry, (long long)-a.ref.yorigin), (long long)(a.ref.alloc_h - a.ref.yorigin - IMP_B));
  j.bsize = R1C_BLOCK_8X8;
  a.jobs[i] = j;
}

__global__ void imp_shares_kernel(ImpArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.w * a.h) return;
  const int y = i / a.w, x = i - y * a.w;
  const r1c_mv mv = a.mvs[(long long)(2 * y) * a.mv_cols + 2 * x];
  const long long reference_x = (long long)x * B_MV + mv.col, reference_y = (long long)y * B_MV + mv.row;
  const float inter_cost = (float)a.inter[i], intra_cost = (float)a.intra[i], future = a.imp[i];
  // (1. - inter_cost / intra_cost).max(0.): f32::max returns the non-NaN operand, like fmaxf
  const float fraction = fmaxf(__fsub_rn(1.f, __fdiv_rn(inter_cost, intra_cost)), 0.f);
  const float amount = __fdiv_rn(__fmul_rn(__fadd_rn(intra_cost, future), fraction), (float)a.n_refs);
  const long long tlx = tdiv(reference_x - (reference_x < 0 ? B_MV - 1 : 0), B_MV) * B_MV;
  const long long tly = tdiv(reference_y - (reference_y < 0 ? B_MV - 1 : 0), B_MV) * B_MV;
  const long long trx = tlx + B_MV, bly = tly + B_MV;
  const long long fx[2] = {trx - reference_x, reference_x + B_MV - trx};
  const long long fy[2] = {bly - reference_y, reference_y + B_MV - bly};
#pragma unroll
  for (int k = 0; k < 4; k++) {  // TL, TR, BL, BR: the order of the reference's propagate() calls
    const long long bx = tdiv(k & 1 ? trx : tlx, B_MV), by = tdiv(k & 2 ? bly : tly, B_MV);
    const float share = __fdiv_rn((float)(fx[k & 1] * fy[k >> 1]), (float)(B_MV * B_MV));
    const bool in = bx >= 0 && by >= 0 && bx < a.w && by < a.h;
    a.keys[4 * i + k] = in ? (uint32_t)(by * a.w + bx) : NO_TARGET;
    a.vals[4 * i + k] = __fmul_rn(amount, share);
  }
}

// sorted (key, value) pairs: the thread at the head of a run adds the run to its target, in order
__global__ void imp_accumulate_kernel(const uint32_t *keys, const float *vals, int n, float *ref_imp) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t key = keys[j];
  if (key == NO_TARGET || (j > 0 && keys[j - 1] == key)) return;
  float acc = ref_imp[key];
  for (int t = j; t < n && keys[t] == key; t++) acc = __fadd_rn(acc, vals[t]);
  ref_imp[key] = acc;
}

__global__ void imp_finalize_kernel(const uint32_t *intra, float *imp, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float intra_cost = (float)intra[i];
  imp[i] = intra_cost > 0.f ? (float)log2((double)__fadd_rn(1.f, __fdiv_rn(imp[i], intra_cost))) : 0.f;
}

}  // namespace

extern "C" int r1c_importance_propagate_dev(r1c_ctx *ctx, const r1c_plane *frame, const r1c_plane *ref,
                                            const r1c_mv *d_mvs, int mv_cols, int mv_rows, const uint32_t *d_intra_costs,
                                            const float *d_importances, int n_refs, int w_in_imp_b, int h_in_imp_b,
                                            int bit_depth, float *d_ref_importances) {
  if (!ctx || !frame || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, frame->px_bytes == ref->px_bytes && bit_depth_ok(bit_depth, frame->px_bytes), R1C_ERR_ARG,
              "bit_depth %d does not match the planes", bit_depth);
  R1C_REQUIRE(ctx, w_in_imp_b >= 0 && h_in_imp_b >= 0 && (long long)w_in_imp_b * h_in_imp_b < (1ll << 28), R1C_ERR_ARG,
              "bad importance-block grid");
  R1C_REQUIRE(ctx, n_refs >= 1 && n_refs <= 3, R1C_ERR_ARG, "n_refs %d (unique references of a frame: 1..3)", n_refs);
  const int n = w_in_imp_b * h_in_imp_b;
  if (n == 0) return R1C_OK;
  R1C_REQUIRE(ctx, d_mvs && d_intra_costs && d_importances && d_ref_importances, R1C_ERR_ARG, "null array");
  R1C_REQUIRE(ctx, mv_cols >= 2 * w_in_imp_b - 1 && mv_rows >= 2 * h_in_imp_b - 1, R1C_ERR_ARG,
              "MV field %dx%d smaller than the importance grid", mv_cols, mv_rows);
  R1C_REQUIRE(ctx, region_in_plane(frame->cfg, 0, 0, (int64_t)w_in_imp_b * IMP_B, (int64_t)h_in_imp_b * IMP_B), R1C_ERR_BOUNDS,
              "the importance-block grid leaves the allocated plane");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *d_jobs, *d_inter, *d_kv, *d_kv2, *d_tmp;
  int rc;
  const size_t pairs = (size_t)4 * n;
  if ((rc = r1c_scratch_dev(ctx, 5, (size_t)n * sizeof(r1c_dist_job), &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 6, (size_t)n * 4, &d_inter))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 7, pairs * 8, &d_kv))) return rc;   // keys | values
  if ((rc = r1c_scratch_dev(ctx, 8, pairs * 8, &d_kv2))) return rc;  // sorted keys | values
  ImpArgs a;
  a.ref = make_view(ref);
  a.mvs = d_mvs;
  a.mv_cols = mv_cols;
  a.intra = d_intra_costs;
  a.inter = (const uint32_t *)d_inter;
  a.imp = d_importances;
  a.n_refs = n_refs;
  a.w = w_in_imp_b;
  a.h = h_in_imp_b;
  a.jobs = (r1c_dist_job *)d_jobs;
  a.keys = (uint32_t *)d_kv;
  a.vals = (float *)((uint32_t *)d_kv + pairs);
  const unsigned grid = (unsigned)((n + 255) / 256);
  imp_jobs_kernel<<<grid, 256, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  if ((rc = r1c_dist_batch_dev(ctx, 1, frame, ref, (const r1c_dist_job *)d_jobs, n, bit_depth, (uint32_t *)d_inter))) return rc;
  imp_shares_kernel<<<grid, 256, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  uint32_t *k2 = (uint32_t *)d_kv2;
  float *v2 = (float *)(k2 + pairs);
  size_t tmp_bytes = 0;
  R1C_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, a.keys, k2, a.vals, v2, (int)pairs, 0, 31, ctx->stream));
  if ((rc = r1c_scratch_dev(ctx, 9, tmp_bytes + 16, &d_tmp))) return rc;
  R1C_CUDA(ctx, cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, a.keys, k2, a.vals, v2, (int)pairs, 0, 31, ctx->stream));
  ctx->launches += 4;  // cub's radix sort passes (histogram + onesweep for 31 key bits)
  imp_accumulate_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, ctx->stream>>>(k2, v2, (int)pairs, d_ref_importances);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_importance_finalize_dev(r1c_ctx *ctx, const uint32_t *d_intra_costs, float *d_importances, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_intra_costs && d_importances)), R1C_ERR_ARG, "null array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  imp_finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_intra_costs, d_importances, n);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_importance_propagate(r1c_ctx *ctx, const r1c_plane *frame, const r1c_plane *ref, const r1c_mv *mvs,
                                        int mv_cols, int mv_rows, const uint32_t *intra_costs, const float *importances,
                                        int n_refs, int w_in_imp_b, int h_in_imp_b, int bit_depth, float *ref_importances) {
  if (!ctx || !frame || !ref) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, w_in_imp_b >= 0 && h_in_imp_b >= 0 && (long long)w_in_imp_b * h_in_imp_b < (1ll << 28), R1C_ERR_ARG,
              "bad importance-block grid");
  const int n = w_in_imp_b * h_in_imp_b;
  if (n == 0) return R1C_OK;
  R1C_REQUIRE(ctx, mvs && intra_costs && importances && ref_importances, R1C_ERR_ARG, "null array");
  R1C_REQUIRE(ctx, mv_cols >= 2 * w_in_imp_b - 1 && mv_rows >= 2 * h_in_imp_b - 1, R1C_ERR_ARG,
              "MV field %dx%d smaller than the importance grid", mv_cols, mv_rows);
  // get_satd asserts that both regions hold the block (src/dist.rs:54-69): the displaced block must lie in the plane
  for (int y = 0; y < h_in_imp_b; y++)
    for (int x = 0; x < w_in_imp_b; x++) {
      const r1c_mv mv = mvs[(size_t)(2 * y) * mv_cols + 2 * x];
      const int64_t rx = ((int64_t)x * B_MV + mv.col) / 8, ry = ((int64_t)y * B_MV + mv.row) / 8;
      R1C_REQUIRE(ctx, region_in_plane(ref->cfg, rx, ry, IMP_B, IMP_B), R1C_ERR_BOUNDS,
                  "importance block (%d, %d): its MV leaves the allocated reference plane", x, y);
    }
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t mb = (size_t)mv_cols * mv_rows * sizeof(r1c_mv), nb = (size_t)n * 4;
  void *d_mv, *d_in, *d_imp, *d_ref, *h_ref;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 0, mb, &d_mv))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, nb, &d_in))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, nb, &d_imp))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 3, nb, &d_ref))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, nb, &h_ref))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_mv, mvs, mb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_in, intra_costs, nb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_imp, importances, nb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_ref, ref_importances, nb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_importance_propagate_dev(ctx, frame, ref, (const r1c_mv *)d_mv, mv_cols, mv_rows, (const uint32_t *)d_in,
                                         (const float *)d_imp, n_refs, w_in_imp_b, h_in_imp_b, bit_depth, (float *)d_ref)))
    return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_ref, d_ref, nb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(ref_importances, h_ref, nb);
  return R1C_OK;
}

extern "C" int r1c_importance_finalize(r1c_ctx *ctx, const uint32_t *intra_costs, float *importances, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (intra_costs && importances)), R1C_ERR_ARG, "null array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t nb = (size_t)n * 4;
  void *d_in, *d_imp, *h_imp;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 1, nb, &d_in))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, nb, &d_imp))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, nb, &h_imp))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_in, intra_costs, nb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_imp, importances, nb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_importance_finalize_dev(ctx, (const uint32_t *)d_in, (float *)d_imp, n))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_imp, d_imp, nb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(importances, h_imp, nb);
  return R1C_OK;
}
