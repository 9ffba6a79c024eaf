// mv_field.cu -- the motion-vector field the searches feed and read:
//   FrameMotionVectors       src/me.rs:32-62     cols x rows MotionVector in 4x4 (MI) units
//   save_block_motion        src/encoder.rs:1432-1444  a block's MV written to every MI cell it covers
//   get_subset_predictors    src/me.rs:82-174    zero MV, coarse MVs quantised to full-pel, EPZS subsets A/B
//                                               (left, top, top-right of the current field + their median)
//                                               and C (left, top, right, bottom, co-located of the previous
//                                               frame's field); zero vectors are dropped except the first entry
// The field is a plain device array of r1c_mv (row-major, `cols` per row), owned by the caller.
// One thread per block: both functions are gathers / small scatters, bound by launch latency.
#include "common.cuh"

namespace {

struct FieldArgs {
  r1c_mv *field;            // current frame (tile == frame): tile_mvs
  const r1c_mv *prev;       // previous frame's field for this reference (nullable): frame_ref.frame_mvs[ref]
  int cols, rows;
  const r1c_block_job *jobs;
  long long n;
  const r1c_mv *mvs;        // save: one MV per job; predictors: 7 coarse MVs per job
  const uint8_t *n_cmvs;    // predictors: number of coarse MVs of each job (0..7)
  r1c_mv *out;              // predictors: 17 per job
  int32_t *out_n;
};

__device__ __forceinline__ bool is_zero(r1c_mv m) { return m.row == 0 && m.col == 0; }
__device__ __forceinline__ r1c_mv fullpel(r1c_mv m) {  // quantize_to_fullpel: i16 division truncates toward zero
  r1c_mv q;
  q.row = (int16_t)((m.row / 8) * 8);
  q.col = (int16_t)((m.col / 8) * 8);
  return q;
}

__global__ void save_motion_kernel(FieldArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const r1c_block_job j = a.jobs[i];
  // a job outside the field (device lists are not validated by the host) writes nothing
  if (j.bsize < 0 || j.bsize >= R1C_BLOCK_SIZES_ALL || j.bo_x < 0 || j.bo_y < 0 || j.bo_x >= a.cols || j.bo_y >= a.rows) return;
  const int x1 = min(j.bo_x + blk_w_of(j.bsize) / 4, a.cols), y1 = min(j.bo_y + blk_h_of(j.bsize) / 4, a.rows);
  const r1c_mv mv = a.mvs[i];
  for (int y = j.bo_y; y < y1; y++)
    for (int x = j.bo_x; x < x1; x++) a.field[(long long)y * a.cols + x] = mv;
}

__global__ void subset_predictors_kernel(FieldArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const r1c_block_job j = a.jobs[i];
  r1c_mv *p = a.out + i * 17;
  int n = 0;
  if (j.bo_x < 0 || j.bo_y < 0 || j.bo_x >= a.cols || j.bo_y >= a.rows) {  // outside the field: no predictors, count 0
    for (int k = 0; k < 17; k++) p[k] = r1c_mv{0, 0};
    a.out_n[i] = 0;
    return;
  }
  p[n++] = r1c_mv{0, 0};
  const int nc = min((int)a.n_cmvs[i], 7);
  for (int k = 0; k < nc; k++) p[n++] = fullpel(a.mvs[i * 7 + k]);
  // EPZS subsets A and B
  int sr = 0, sc = 0, nm = 0;
  const int x = j.bo_x, y = j.bo_y;
  if (x > 0) {
    const r1c_mv left = a.field[(long long)y * a.cols + x - 1];
    sr += left.row, sc += left.col, nm++;
    if (!is_zero(left)) p[n++] = left;
  }
  if (y > 0) {
    const r1c_mv top = a.field[(long long)(y - 1) * a.cols + x];
    sr += top.row, sc += top.col, nm++;
    if (!is_zero(top)) p[n++] = top;
    if (x < a.cols - 1) {
      const r1c_mv tr = a.field[(long long)(y - 1) * a.cols + x + 1];
      sr += tr.row, sc += tr.col, nm++;
      if (!is_zero(tr)) p[n++] = tr;
    }
  }
  if (nm) {
    r1c_mv med;  // i16 sums (wrapping like the release build), i16 division truncating toward zero
    med.row = (int16_t)((int16_t)sr / (int16_t)nm);
    med.col = (int16_t)((int16_t)sc / (int16_t)nm);
    med = fullpel(med);
    if (!is_zero(med)) p[n++] = med;
  }
  // EPZS subset C
  if (a.prev) {
    if (x > 0) {
      const r1c_mv m = a.prev[(long long)y * a.cols + x - 1];
      if (!is_zero(m)) p[n++] = m;
    }
    if (y > 0) {
      const r1c_mv m = a.prev[(long long)(y - 1) * a.cols + x];
      if (!is_zero(m)) p[n++] = m;
    }
    if (x < a.cols - 1) {
      const r1c_mv m = a.prev[(long long)y * a.cols + x + 1];
      if (!is_zero(m)) p[n++] = m;
    }
    if (y < a.rows - 1) {
      const r1c_mv m = a.prev[(long long)(y + 1) * a.cols + x];
      if (!is_zero(m)) p[n++] = m;
    }
    const r1c_mv m = a.prev[(long long)y * a.cols + x];
    if (!is_zero(m)) p[n++] = m;
  }
  a.out_n[i] = n;
  for (int k = n; k < 17; k++) p[k] = p[n - 1];  // padding by repetition (what r1c_diamond_search_batch expects)
}

int check_field(r1c_ctx *ctx, int cols, int rows, int64_t n) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, cols > 0 && rows > 0 && n >= 0 && n < (1ll << 28), R1C_ERR_ARG, "bad field geometry or job count");
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_save_block_motion_batch_dev(r1c_ctx *ctx, r1c_mv *d_field, int cols, int rows, const r1c_block_job *d_jobs,
                                               const r1c_mv *d_mvs, int64_t n) {
  int rc = check_field(ctx, cols, rows, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (d_field && d_jobs && d_mvs), R1C_ERR_ARG, "null field/job array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  FieldArgs a = {};
  a.field = d_field;
  a.cols = cols;
  a.rows = rows;
  a.jobs = d_jobs;
  a.n = n;
  a.mvs = d_mvs;
  save_motion_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_subset_predictors_batch_dev(r1c_ctx *ctx, const r1c_mv *d_field, const r1c_mv *d_prev_field, int cols, int rows,
                                               const r1c_block_job *d_jobs, int64_t n, const r1c_mv *d_cmvs, const uint8_t *d_n_cmvs,
                                               r1c_mv *d_predictors, int32_t *d_counts) {
  int rc = check_field(ctx, cols, rows, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (d_field && d_jobs && d_cmvs && d_n_cmvs && d_predictors && d_counts), R1C_ERR_ARG, "null array");
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  FieldArgs a = {};
  a.field = (r1c_mv *)d_field;
  a.prev = d_prev_field;
  a.cols = cols;
  a.rows = rows;
  a.jobs = d_jobs;
  a.n = n;
  a.mvs = d_cmvs;
  a.n_cmvs = d_n_cmvs;
  a.out = d_predictors;
  a.out_n = d_counts;
  subset_predictors_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// host-array forms: field(s) in / out through staging copies
extern "C" int r1c_save_block_motion_batch(r1c_ctx *ctx, r1c_mv *field, int cols, int rows, const r1c_block_job *jobs,
                                           const r1c_mv *mvs, int64_t n) {
  int rc = check_field(ctx, cols, rows, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (field && jobs && mvs), R1C_ERR_ARG, "null field/job array");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, jobs[i].bsize >= 0 && jobs[i].bsize < R1C_BLOCK_SIZES_ALL && jobs[i].bo_x >= 0 && jobs[i].bo_y >= 0 &&
                         jobs[i].bo_x < cols && jobs[i].bo_y < rows,
                R1C_ERR_ARG, "block job %lld outside the field", (long long)i);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t fb = (size_t)cols * rows * sizeof(r1c_mv), jb = (size_t)n * sizeof(r1c_block_job), mb = (size_t)n * sizeof(r1c_mv);
  void *d_f, *d_j, *d_m;
  if ((rc = r1c_scratch_dev(ctx, 0, fb, &d_f))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, jb, &d_j))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, mb, &d_m))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_f, field, fb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_j, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_m, mvs, mb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_save_block_motion_batch_dev(ctx, (r1c_mv *)d_f, cols, rows, (const r1c_block_job *)d_j, (const r1c_mv *)d_m, n))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(field, d_f, fb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_subset_predictors_batch(r1c_ctx *ctx, const r1c_mv *field, const r1c_mv *prev_field, int cols, int rows,
                                           const r1c_block_job *jobs, int64_t n, const r1c_mv *cmvs, const uint8_t *n_cmvs,
                                           r1c_mv *predictors, int32_t *counts) {
  int rc = check_field(ctx, cols, rows, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (field && jobs && cmvs && n_cmvs && predictors && counts), R1C_ERR_ARG, "null array");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++)
    R1C_REQUIRE(ctx, jobs[i].bo_x >= 0 && jobs[i].bo_y >= 0 && jobs[i].bo_x < cols && jobs[i].bo_y < rows && n_cmvs[i] <= 7, R1C_ERR_ARG,
                "block job %lld outside the field or more than 7 coarse MVs", (long long)i);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t fb = (size_t)cols * rows * sizeof(r1c_mv), jb = (size_t)n * sizeof(r1c_block_job), cb = (size_t)n * 7 * sizeof(r1c_mv);
  const size_t pb = (size_t)n * 17 * sizeof(r1c_mv), nb = (size_t)n * sizeof(int32_t);
  void *d_f, *d_p = nullptr, *d_j, *d_c, *d_nc, *d_o, *d_on;
  if ((rc = r1c_scratch_dev(ctx, 0, fb, &d_f))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, jb, &d_j))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 2, cb, &d_c))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 5, (size_t)n, &d_nc))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 6, pb, &d_o))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 7, nb, &d_on))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_f, field, fb, cudaMemcpyHostToDevice, ctx->stream));
  if (prev_field) {
    if ((rc = r1c_scratch_dev(ctx, 8, fb, &d_p))) return rc;
    R1C_CUDA(ctx, cudaMemcpyAsync(d_p, prev_field, fb, cudaMemcpyHostToDevice, ctx->stream));
  }
  R1C_CUDA(ctx, cudaMemcpyAsync(d_j, jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_c, cmvs, cb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_nc, n_cmvs, (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = r1c_subset_predictors_batch_dev(ctx, (const r1c_mv *)d_f, (const r1c_mv *)d_p, cols, rows, (const r1c_block_job *)d_j, n,
                                            (const r1c_mv *)d_c, (const uint8_t *)d_nc, (r1c_mv *)d_o, (int32_t *)d_on)))
    return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(predictors, d_o, pb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(counts, d_on, nb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}
