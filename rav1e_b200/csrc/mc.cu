// mc.cu -- batched sub-pel motion compensation: put_8tap / prep_8tap / mc_avg
// (reference src/mc.rs:410-706; semantics of put_8tap_ref :213-307,
// prep_8tap_ref :310-387, mc_avg_ref :389-408, get_filter :201-210).
//
// Work decomposition.  A job (one block, 2..128 px on a side) is cut into
// ITEMS of 4 columns x up to 32 rows; one thread owns one item and walks down
// its rows with the 8-row window of horizontally filtered values held in
// registers, so the separable filter needs no shared memory: per row a thread
// loads the 11 source pixels of its strip as aligned 32-bit words (neighbouring
// strips overlap in L1), does 4 x 8 horizontal MACs, then 4 x 8 vertical MACs.
// Item counts per job go through a prefix sum (cub) so that job lists mixing
// every block size still fill all lanes; the main kernel is persistent (grid =
// a multiple of the SM count) and reads the item total from device memory, so
// the `_dev` entry points never synchronise with the host.
//
// Arithmetic follows the `_ref` functions literally: i32 accumulation, the
// horizontal intermediate truncated to i16 (`as i16`, src/mc.rs:281,366),
// intermediate_bits = 4 (8/10-bit) or 2 (12-bit), and the four (col_frac,
// row_frac) cases with their own rounding.
// Algorithmic bytes: put/prep 2*px_bytes per output pixel (1 read + 1 write,
// halo served by L1/L2); mc_avg 2*2 B read + px_bytes written.
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace {

// SUBPEL_FILTERS, src/mc.rs:70-179: [bank][frac][tap]; banks 0..3 REGULAR, SMOOTH,
// SHARP, BILINEAR; 4 / 5 = the 4-tap REGULAR / SMOOTH(+SHARP) variants for
// lengths <= 4.  frac 0 (unit tap 128) never reaches a filter in the kernels.
__constant__ short c_filters[6][16][8] = {
    {{0, 0, 0, 128, 0, 0, 0, 0}, {0, 2, -6, 126, 8, -2, 0, 0}, {0, 2, -10, 122, 18, -4, 0, 0}, {0, 2, -12, 116, 28, -8, 2, 0},
     {0, 2, -14, 110, 38, -10, 2, 0}, {0, 2, -14, 102, 48, -12, 2, 0}, {0, 2, -16, 94, 58, -12, 2, 0}, {0, 2, -14, 84, 66, -12, 2, 0},
     {0, 2, -14, 76, 76, -14, 2, 0}, {0, 2, -12, 66, 84, -14, 2, 0}, {0, 2, -12, 58, 94, -16, 2, 0}, {0, 2, -12, 48, 102, -14, 2, 0},
     {0, 2, -10, 38, 110, -14, 2, 0}, {0, 2, -8, 28, 116, -12, 2, 0}, {0, 0, -4, 18, 122, -10, 2, 0}, {0, 0, -2, 8, 126, -6, 2, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0}, {0, 2, 28, 62, 34, 2, 0, 0}, {0, 0, 26, 62, 36, 4, 0, 0}, {0, 0, 22, 62, 40, 4, 0, 0},
     {0, 0, 20, 60, 42, 6, 0, 0}, {0, 0, 18, 58, 44, 8, 0, 0}, {0, 0, 16, 56, 46, 10, 0, 0}, {0, -2, 16, 54, 48, 12, 0, 0},
     {0, -2, 14, 52, 52, 14, -2, 0}, {0, 0, 12, 48, 54, 16, -2, 0}, {0, 0, 10, 46, 56, 16, 0, 0}, {0, 0, 8, 44, 58, 18, 0, 0},
     {0, 0, 6, 42, 60, 20, 0, 0}, {0, 0, 4, 40, 62, 22, 0, 0}, {0, 0, 4, 36, 62, 26, 0, 0}, {0, 0, 2, 34, 62, 28, 2, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0}, {-2, 2, -6, 126, 8, -2, 2, 0}, {-2, 6, -12, 124, 16, -6, 4, -2}, {-2, 8, -18, 120, 26, -10, 6, -2},
     {-4, 10, -22, 116, 38, -14, 6, -2}, {-4, 10, -22, 108, 48, -18, 8, -2}, {-4, 10, -24, 100, 60, -20, 8, -2}, {-4, 10, -24, 90, 70, -22, 10, -2},
     {-4, 12, -24, 80, 80, -24, 12, -4}, {-2, 10, -22, 70, 90, -24, 10, -4}, {-2, 8, -20, 60, 100, -24, 10, -4}, {-2, 8, -18, 48, 108, -22, 10, -4},
     {-2, 6, -14, 38, 116, -22, 10, -4}, {-2, 6, -10, 26, 120, -18, 8, -2}, {-2, 4, -6, 16, 124, -12, 6, -2}, {0, 2, -2, 8, 126, -6, 2, -2}},
    {{0, 0, 0, 128, 0, 0, 0, 0}, {0, 0, 0, 120, 8, 0, 0, 0}, {0, 0, 0, 112, 16, 0, 0, 0}, {0, 0, 0, 104, 24, 0, 0, 0},
     {0, 0, 0, 96, 32, 0, 0, 0}, {0, 0, 0, 88, 40, 0, 0, 0}, {0, 0, 0, 80, 48, 0, 0, 0}, {0, 0, 0, 72, 56, 0, 0, 0},
     {0, 0, 0, 64, 64, 0, 0, 0}, {0, 0, 0, 56, 72, 0, 0, 0}, {0, 0, 0, 48, 80, 0, 0, 0}, {0, 0, 0, 40, 88, 0, 0, 0},
     {0, 0, 0, 32, 96, 0, 0, 0}, {0, 0, 0, 24, 104, 0, 0, 0}, {0, 0, 0, 16, 112, 0, 0, 0}, {0, 0, 0, 8, 120, 0, 0, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0}, {0, 0, -4, 126, 8, -2, 0, 0}, {0, 0, -8, 122, 18, -4, 0, 0}, {0, 0, -10, 116, 28, -6, 0, 0},
     {0, 0, -12, 110, 38, -8, 0, 0}, {0, 0, -12, 102, 48, -10, 0, 0}, {0, 0, -14, 94, 58, -10, 0, 0}, {0, 0, -12, 84, 66, -10, 0, 0},
     {0, 0, -12, 76, 76, -12, 0, 0}, {0, 0, -10, 66, 84, -12, 0, 0}, {0, 0, -10, 58, 94, -14, 0, 0}, {0, 0, -10, 48, 102, -12, 0, 0},
     {0, 0, -8, 38, 110, -12, 0, 0}, {0, 0, -6, 28, 116, -10, 0, 0}, {0, 0, -4, 18, 122, -8, 0, 0}, {0, 0, -2, 8, 126, -4, 0, 0}},
    {{0, 0, 0, 128, 0, 0, 0, 0}, {0, 0, 30, 62, 34, 2, 0, 0}, {0, 0, 26, 62, 36, 4, 0, 0}, {0, 0, 22, 62, 40, 4, 0, 0},
     {0, 0, 20, 60, 42, 6, 0, 0}, {0, 0, 18, 58, 44, 8, 0, 0}, {0, 0, 16, 56, 46, 10, 0, 0}, {0, 0, 14, 54, 48, 12, 0, 0},
     {0, 0, 12, 52, 52, 12, 0, 0}, {0, 0, 12, 48, 54, 14, 0, 0}, {0, 0, 10, 46, 56, 16, 0, 0}, {0, 0, 8, 44, 58, 18, 0, 0},
     {0, 0, 6, 42, 60, 20, 0, 0}, {0, 0, 4, 40, 62, 22, 0, 0}, {0, 0, 4, 36, 62, 26, 0, 0}, {0, 0, 2, 34, 62, 30, 0, 0}}};

constexpr int SEG_ROWS = 32;  // rows per item
constexpr int MC_THREADS = 128;

__host__ __device__ inline int items_of(int w, int h) { return ((w + 3) >> 2) * ((h + SEG_ROWS - 1) / SEG_ROWS); }

// get_filter, src/mc.rs:201-210
__device__ __forceinline__ int filter_bank(int mode, int length) {
  return (mode == R1C_BILINEAR || length > 4) ? mode : ((mode < 1 ? mode : 1) + 4);
}

__device__ __forceinline__ int rshift_round(int v, int bit) { return (v + ((1 << bit) >> 1)) >> bit; }

template <typename JOB>
__global__ void count_items_kernel(const JOB *jobs, long long n, long long *counts) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) counts[i] = items_of(jobs[i].w, jobs[i].h);
  if (i == n) counts[i] = 0;
}

// largest j in [0, n) with offs[j] <= item (offs is the exclusive prefix sum, offs[n] = total)
__device__ __forceinline__ long long find_job(const long long *offs, long long n, long long item) {
  long long lo = 0, hi = n;  // invariant: offs[lo] <= item < offs[hi]
  while (hi - lo > 1) {
    long long mid = (lo + hi) >> 1;
    if (__ldg(offs + mid) <= item) lo = mid;
    else hi = mid;
  }
  return lo;
}

// Source pixels of one row of a strip: px[k] = pixel at column x0 - 3 + k.
// HAS_H: k = 0 .. ncols+6 are needed (horizontal filter footprint); otherwise
// only k = 3 .. 3+ncols-1 (the strip itself).  Only aligned 32-bit words that
// contain a needed pixel are touched, so the load never leaves the allocation.
template <int PB, bool HAS_H>
__device__ __forceinline__ void load_strip_row(const uint8_t *row, int x0, int ncols, int (&px)[11]) {
  constexpr int KLO = HAS_H ? 0 : 3;
  const int nneed = HAS_H ? ncols + 7 : ncols;
  const uint8_t *first = row + (long long)(x0 - 3 + KLO) * PB;
  const uintptr_t a0 = (uintptr_t)first & ~(uintptr_t)3;
  const uint32_t *wp = (const uint32_t *)a0;
  const int nwords = (int)(((uintptr_t)first + (uintptr_t)(nneed * PB) - a0 + 3) >> 2);
  constexpr int NA = HAS_H ? (PB == 1 ? 3 : 6) : (PB == 1 ? 1 : 2);  // realigned words
  uint32_t w[NA + 1];
#pragma unroll
  for (int i = 0; i < NA + 1; i++) w[i] = i < nwords ? __ldg(wp + i) : 0u;
  const uint32_t sh = (uint32_t)((uintptr_t)first & 3) * 8;
  uint32_t a[NA];
#pragma unroll
  for (int i = 0; i < NA; i++) a[i] = __funnelshift_r(w[i], w[i + 1], sh);
#pragma unroll
  for (int k = 0; k < 11; k++) {
    constexpr int PPW = 4 / PB;  // pixels per word
    const int q = k - KLO;
    if (q >= 0 && q < NA * PPW) px[k] = (int)((a[q / PPW] >> ((q % PPW) * 8 * PB)) & (PB == 1 ? 0xffu : 0xffffu));
    else px[k] = 0;
  }
}

struct McArgs {
  PlaneView src;
  uint8_t *dst0;  // put: address of dst plane coordinate (0,0)
  int dst_stride;
  int16_t *tmp;   // prep
  const r1c_mc_job *jobs;
  const long long *offs;
  long long n;
  int bit_depth;
};

// KIND 0 = put (clamped pixels into the dst plane), 1 = prep (i16, dense w*h at tmp_off),
// 2 = put into a dense w*h pixel block at tmp_off (pixels; used by the sub-pel search)
template <int PB, int KIND>
__global__ void __launch_bounds__(MC_THREADS) mc_kernel(McArgs a) {
  const long long total = a.offs[a.n];
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;
  for (long long item = (long long)blockIdx.x * MC_THREADS + threadIdx.x; item < total;
       item += (long long)gridDim.x * MC_THREADS) {
    const long long ji = find_job(a.offs, a.n, item);
    const r1c_mc_job j = a.jobs[ji];
    const int local = (int)(item - a.offs[ji]);
    const int nstrips = (j.w + 3) >> 2;
    const int seg = local / nstrips, strip = local - seg * nstrips;
    const int x0 = strip * 4;
    const int ncols = min(4, j.w - x0);
    const int r0 = seg * SEG_ROWS, r1 = min((int)j.h, r0 + SEG_ROWS);
    const bool has_h = j.col_frac != 0, has_v = j.row_frac != 0;

    int xf[8], yf[8];
    {
      const short *fx = c_filters[filter_bank(j.mode_x, j.w)][j.col_frac & 15];
      const short *fy = c_filters[filter_bank(j.mode_y, j.h)][j.row_frac & 15];
#pragma unroll
      for (int k = 0; k < 8; k++) {
        xf[k] = fx[k];
        yf[k] = fy[k];
      }
    }
    const uint8_t *src = (const uint8_t *)a.src.p0 + ((long long)j.src_y * a.src.stride + j.src_x) * PB;
    const long long sstride = (long long)a.src.stride * PB;
    const int vshift = has_h ? (KIND == 1 ? 7 : 7 + ib) : (KIND == 1 ? 7 - ib : 7);

    int win[4][8];  // win[c][k]: horizontally processed value of row (current - 7 + k)
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
      for (int k = 0; k < 8; k++) win[c][k] = 0;

    const int row_first = has_v ? r0 - 3 : r0;
    const int row_last = has_v ? r1 + 4 : r1;  // exclusive
    for (int rb = row_first; rb < row_last; rb += 8) {
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int r = rb + u;  // source row relative to the block origin
        if (r < row_last) {
          int px[11];
          if (has_h) load_strip_row<PB, true>(src + r * sstride, x0, ncols, px);
          else load_strip_row<PB, false>(src + r * sstride, x0, ncols, px);
          int hv[4];  // value entering the vertical stage (or the output when !has_v)
#pragma unroll
          for (int c = 0; c < 4; c++) {
            if (has_h) {
              int s = 0;
#pragma unroll
              for (int k = 0; k < 8; k++) s += xf[k] * px[c + k];
              hv[c] = rshift_round(s, 7 - ib);  // i32; cast to i16 where the reference does
            } else {
              hv[c] = px[c + 3];
            }
          }
          int outv[4];
          bool emit = false;
          int orow = 0;
          if (has_v) {
            // slot (u) of the rotating window receives this row; the window is
            // complete once 8 rows are in: rows r-7 .. r produce output row r-4
#pragma unroll
            for (int c = 0; c < 4; c++) win[c][u] = has_h ? (int)(int16_t)hv[c] : hv[c];
            orow = r - 4;
            if (orow >= r0) {
              emit = true;
#pragma unroll
              for (int c = 0; c < 4; c++) {
                int s = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) s += yf[k] * win[c][(u + 1 + k) & 7];
                s = rshift_round(s, vshift);
                outv[c] = KIND == 1 ? (int)(int16_t)s : min(max(s, 0), maxv);
              }
            }
          } else {
            emit = true;
            orow = r;
#pragma unroll
            for (int c = 0; c < 4; c++) {
              if (has_h) outv[c] = KIND == 1 ? (int)(int16_t)hv[c] : min(max(rshift_round(hv[c], ib), 0), maxv);
              else outv[c] = KIND == 1 ? (int)(int16_t)((int16_t)hv[c] << ib) : hv[c];
            }
          }
          if (emit) {
            if (KIND == 0 || KIND == 2) {
              uint8_t *d = KIND == 0 ? a.dst0 + ((long long)(j.dst_y + orow) * a.dst_stride + j.dst_x + x0) * PB
                                     : (uint8_t *)a.tmp + ((long long)j.tmp_off + (long long)orow * j.w + x0) * PB;
              if (PB == 1) {
                if (ncols == 4 && (((uintptr_t)d) & 3) == 0) {
                  *(uint32_t *)d = (uint32_t)outv[0] | ((uint32_t)outv[1] << 8) | ((uint32_t)outv[2] << 16) | ((uint32_t)outv[3] << 24);
                } else {
#pragma unroll
                  for (int c = 0; c < 4; c++)
                    if (c < ncols) d[c] = (uint8_t)outv[c];
                }
              } else {
                uint16_t *d16 = (uint16_t *)d;
                if (ncols == 4 && (((uintptr_t)d) & 7) == 0) {
                  uint2 v;
                  v.x = (uint32_t)outv[0] | ((uint32_t)outv[1] << 16);
                  v.y = (uint32_t)outv[2] | ((uint32_t)outv[3] << 16);
                  *(uint2 *)d = v;
                } else {
#pragma unroll
                  for (int c = 0; c < 4; c++)
                    if (c < ncols) d16[c] = (uint16_t)outv[c];
                }
              }
            } else {
              int16_t *t = a.tmp + j.tmp_off + (long long)orow * j.w + x0;
              if (ncols == 4 && (((uintptr_t)t) & 7) == 0) {
                uint2 v;
                v.x = ((uint32_t)outv[0] & 0xffff) | ((uint32_t)outv[1] << 16);
                v.y = ((uint32_t)outv[2] & 0xffff) | ((uint32_t)outv[3] << 16);
                *(uint2 *)t = v;
              } else {
#pragma unroll
                for (int c = 0; c < 4; c++)
                  if (c < ncols) t[c] = (int16_t)outv[c];
              }
            }
          }
        }
      }
    }
  }
}

struct AvgArgs {
  uint8_t *dst0;
  int dst_stride;
  const int16_t *t1, *t2;
  const r1c_avg_job *jobs;
  const long long *offs;
  long long n;
  int bit_depth;
};

// mc_avg_ref, src/mc.rs:389-408
template <int PB>
__global__ void __launch_bounds__(MC_THREADS) avg_kernel(AvgArgs a) {
  const long long total = a.offs[a.n];
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;
  for (long long item = (long long)blockIdx.x * MC_THREADS + threadIdx.x; item < total;
       item += (long long)gridDim.x * MC_THREADS) {
    const long long ji = find_job(a.offs, a.n, item);
    const r1c_avg_job j = a.jobs[ji];
    const int local = (int)(item - a.offs[ji]);
    const int nstrips = (j.w + 3) >> 2;
    const int seg = local / nstrips, strip = local - seg * nstrips;
    const int x0 = strip * 4, ncols = min(4, j.w - x0);
    const int r0 = seg * SEG_ROWS, r1 = min((int)j.h, r0 + SEG_ROWS);
    for (int r = r0; r < r1; r++) {
      const long long o = (long long)j.tmp_off + (long long)r * j.w + x0;
      uint8_t *d = a.dst0 + ((long long)(j.dst_y + r) * a.dst_stride + j.dst_x + x0) * PB;
#pragma unroll
      for (int c = 0; c < 4; c++) {
        if (c < ncols) {
          int v = rshift_round((int)a.t1[o + c] + (int)a.t2[o + c], ib + 1);
          v = min(max(v, 0), maxv);
          if (PB == 1) d[c] = (uint8_t)v;
          else ((uint16_t *)d)[c] = (uint16_t)v;
        }
      }
    }
  }
}

template <typename JOB>
int build_offsets(r1c_ctx *ctx, const JOB *d_jobs, int64_t n, long long **d_offs_out) {
  void *d_counts, *d_offs, *d_tmp;
  int rc;
  const size_t bytes = (size_t)(n + 1) * sizeof(long long);
  if ((rc = r1c_scratch_dev(ctx, 2, bytes, &d_counts))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 3, bytes, &d_offs))) return rc;
  count_items_kernel<JOB><<<(unsigned)((n + 1 + 255) / 256), 256, 0, ctx->stream>>>(d_jobs, n, (long long *)d_counts);
  R1C_LAUNCH_CHECK(ctx);
  size_t tmp_bytes = 0;
  R1C_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (long long *)d_counts, (long long *)d_offs, (int)(n + 1), ctx->stream));
  if ((rc = r1c_scratch_dev(ctx, 4, tmp_bytes + 16, &d_tmp))) return rc;
  R1C_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, (long long *)d_counts, (long long *)d_offs, (int)(n + 1), ctx->stream));
  ctx->launches += 2;  // cub's scan kernels
  *d_offs_out = (long long *)d_offs;
  return R1C_OK;
}

int check_mc(r1c_ctx *ctx, const r1c_plane *src, int bit_depth, int64_t n, const void *jobs) {
  if (!ctx || !src) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, src->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, src->px_bytes);
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 31) - 2 && (n == 0 || jobs), R1C_ERR_ARG, "bad job array");
  return R1C_OK;
}

int grid_for(r1c_ctx *ctx) { return ctx->sm_count * 8; }

int launch_mc(r1c_ctx *ctx, int kind, r1c_plane *dst, int16_t *d_tmp, const r1c_plane *src,
              const r1c_mc_job *d_jobs, int64_t n, int bit_depth) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  long long *d_offs;
  int rc = build_offsets(ctx, d_jobs, n, &d_offs);
  if (rc) return rc;
  McArgs a;
  a.src = make_view(src);
  a.dst0 = nullptr;
  a.dst_stride = 0;
  if (kind == 0) {
    PlaneView dv = make_view(dst);
    a.dst0 = (uint8_t *)dv.p0;
    a.dst_stride = dv.stride;
  }
  a.tmp = d_tmp;
  a.jobs = d_jobs;
  a.offs = d_offs;
  a.n = n;
  a.bit_depth = bit_depth;
  const int grid = grid_for(ctx);
  r1c_prof_begin(ctx, ctx->prof_mc_slot ? ctx->prof_mc_slot : R1C_PROF_MC);
  if (src->px_bytes == 1) {
    if (kind == 0) mc_kernel<1, 0><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
    else if (kind == 1) mc_kernel<1, 1><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
    else mc_kernel<1, 2><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
  } else {
    if (kind == 0) mc_kernel<2, 0><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
    else if (kind == 1) mc_kernel<2, 1><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
    else mc_kernel<2, 2><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
  }
  r1c_prof_end(ctx, ctx->prof_mc_slot ? ctx->prof_mc_slot : R1C_PROF_MC);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// host-side validation of one MC job: the source footprint of each (col_frac,
// row_frac) case must lie inside the allocated plane
int validate_mc_job(r1c_ctx *ctx, const r1c_plane *src, const r1c_mc_job &j, int64_t i) {
  R1C_REQUIRE(ctx, j.w >= 1 && j.w <= 128 && j.h >= 1 && j.h <= 128, R1C_ERR_ARG, "mc job %lld: bad size %dx%d", (long long)i, j.w, j.h);
  R1C_REQUIRE(ctx, j.col_frac >= 0 && j.col_frac < 16 && j.row_frac >= 0 && j.row_frac < 16, R1C_ERR_ARG, "mc job %lld: frac out of range", (long long)i);
  R1C_REQUIRE(ctx, j.mode_x >= 0 && j.mode_x < 4 && j.mode_y >= 0 && j.mode_y < 4, R1C_ERR_ARG, "mc job %lld: bad filter mode", (long long)i);
  const int64_t x = j.src_x - (j.col_frac ? 3 : 0), y = j.src_y - (j.row_frac ? 3 : 0);
  const int64_t w = j.w + (j.col_frac ? 7 : 0), h = j.h + (j.row_frac ? 7 : 0);
  R1C_REQUIRE(ctx, region_in_plane(src->cfg, x, y, w, h), R1C_ERR_BOUNDS, "mc job %lld: source footprint leaves the allocated plane", (long long)i);
  return R1C_OK;
}

}  // namespace

// internal (subpel.cu): put_8tap of every job into a dense pixel buffer, block i at
// d_dense + jobs[i].tmp_off pixels, row stride = jobs[i].w; jobs with w == 0 are skipped
int r1c_mc_put_dense_dev(r1c_ctx *ctx, void *d_dense, const r1c_plane *src, const r1c_mc_job *d_jobs,
                         int64_t n, int bit_depth) {
  return launch_mc(ctx, 2, nullptr, (int16_t *)d_dense, src, d_jobs, n, bit_depth);
}

extern "C" void r1c_mv_params(const r1c_plane_cfg *cfg, int64_t po_x, int64_t po_y, int mv_row, int mv_col,
                              int *row_frac, int *col_frac, int64_t *src_x, int64_t *src_y) {
  // get_params, src/predict.rs:267-284
  const int sr = 3 + (int)cfg->ydec, sc = 3 + (int)cfg->xdec;
  const int ro = mv_row >> sr, co = mv_col >> sc;
  *row_frac = (mv_row - (ro << sr)) << (4 - sr);
  *col_frac = (mv_col - (co << sc)) << (4 - sc);
  int64_t qx = po_x + co - 3, qy = po_y + ro - 3;
  // PlaneSlice::clamp, src/frame/plane.rs:521-533
  if (qx > cfg->width) qx = cfg->width;
  if (qx < -cfg->xorigin) qx = -cfg->xorigin;
  if (qy > cfg->height) qy = cfg->height;
  if (qy < -cfg->yorigin) qy = -cfg->yorigin;
  *src_x = qx + 3;
  *src_y = qy + 3;
}

extern "C" int r1c_put_8tap_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src,
                                      const r1c_mc_job *d_jobs, int64_t n, int bit_depth) {
  int rc = check_mc(ctx, src, bit_depth, n, d_jobs);
  if (rc) return rc;
  R1C_REQUIRE(ctx, dst && dst->px_bytes == src->px_bytes, R1C_ERR_ARG, "dst/src pixel types differ");
  return launch_mc(ctx, 0, dst, nullptr, src, d_jobs, n, bit_depth);
}

extern "C" int r1c_prep_8tap_batch_dev(r1c_ctx *ctx, int16_t *d_tmp, const r1c_plane *src,
                                       const r1c_mc_job *d_jobs, int64_t n, int bit_depth) {
  int rc = check_mc(ctx, src, bit_depth, n, d_jobs);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || d_tmp, R1C_ERR_ARG, "null tmp");
  return launch_mc(ctx, 1, nullptr, d_tmp, src, d_jobs, n, bit_depth);
}

extern "C" int r1c_put_8tap_batch(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src,
                                  const r1c_mc_job *jobs, int64_t n, int bit_depth) {
  int rc = check_mc(ctx, src, bit_depth, n, jobs);
  if (rc) return rc;
  R1C_REQUIRE(ctx, dst && dst->px_bytes == src->px_bytes, R1C_ERR_ARG, "dst/src pixel types differ");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++) {
    if ((rc = validate_mc_job(ctx, src, jobs[i], i))) return rc;
    // dst region assert of put_8tap, src/mc.rs:415-422
    R1C_REQUIRE(ctx, region_in_plane(dst->cfg, jobs[i].dst_x, jobs[i].dst_y, jobs[i].w, jobs[i].h), R1C_ERR_BOUNDS, "mc job %lld: dst block leaves the allocated plane", (long long)i);
  }
  void *h_jobs, *d_jobs;
  const size_t jb = (size_t)n * sizeof(r1c_mc_job);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  memcpy(h_jobs, jobs, jb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_mc(ctx, 0, dst, nullptr, src, (const r1c_mc_job *)d_jobs, n, bit_depth))) return rc;
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_prep_8tap_batch(r1c_ctx *ctx, int16_t *tmp, int64_t tmp_len, const r1c_plane *src,
                                   const r1c_mc_job *jobs, int64_t n, int bit_depth) {
  int rc = check_mc(ctx, src, bit_depth, n, jobs);
  if (rc) return rc;
  if (n == 0) return R1C_OK;
  R1C_REQUIRE(ctx, tmp && tmp_len > 0, R1C_ERR_ARG, "null tmp");
  for (int64_t i = 0; i < n; i++) {
    if ((rc = validate_mc_job(ctx, src, jobs[i], i))) return rc;
    // `tmp.len() >= width * height`, src/mc.rs:525-528
    R1C_REQUIRE(ctx, jobs[i].tmp_off >= 0 && (int64_t)jobs[i].tmp_off + (int64_t)jobs[i].w * jobs[i].h <= tmp_len, R1C_ERR_BOUNDS, "mc job %lld: tmp is smaller than the compute block", (long long)i);
  }
  void *h_jobs, *d_jobs, *h_tmp, *d_tmp;
  const size_t jb = (size_t)n * sizeof(r1c_mc_job), tb = (size_t)tmp_len * 2;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, tb, &h_tmp))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, tb, &d_tmp))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_tmp, tmp, tb);  // elements no job writes keep the caller's values
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_tmp, h_tmp, tb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_mc(ctx, 1, nullptr, (int16_t *)d_tmp, src, (const r1c_mc_job *)d_jobs, n, bit_depth))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_tmp, d_tmp, tb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(tmp, h_tmp, tb);
  return R1C_OK;
}

static int launch_avg(r1c_ctx *ctx, r1c_plane *dst, const int16_t *d_t1, const int16_t *d_t2,
                      const r1c_avg_job *d_jobs, int64_t n, int bit_depth) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  long long *d_offs;
  int rc = build_offsets(ctx, d_jobs, n, &d_offs);
  if (rc) return rc;
  AvgArgs a;
  PlaneView dv = make_view(dst);
  a.dst0 = (uint8_t *)dv.p0;
  a.dst_stride = dv.stride;
  a.t1 = d_t1;
  a.t2 = d_t2;
  a.jobs = d_jobs;
  a.offs = d_offs;
  a.n = n;
  a.bit_depth = bit_depth;
  r1c_prof_begin(ctx, R1C_PROF_MC);
  if (dst->px_bytes == 1) avg_kernel<1><<<grid_for(ctx), MC_THREADS, 0, ctx->stream>>>(a);
  else avg_kernel<2><<<grid_for(ctx), MC_THREADS, 0, ctx->stream>>>(a);
  r1c_prof_end(ctx, R1C_PROF_MC);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

extern "C" int r1c_mc_avg_batch_dev(r1c_ctx *ctx, r1c_plane *dst, const int16_t *d_tmp1, const int16_t *d_tmp2,
                                    const r1c_avg_job *d_jobs, int64_t n, int bit_depth) {
  if (!ctx || !dst) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, dst->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, dst->px_bytes);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (d_jobs && d_tmp1 && d_tmp2)), R1C_ERR_ARG, "null array");
  return launch_avg(ctx, dst, d_tmp1, d_tmp2, d_jobs, n, bit_depth);
}

extern "C" int r1c_mc_avg_batch(r1c_ctx *ctx, r1c_plane *dst, const int16_t *tmp1, const int16_t *tmp2,
                                int64_t tmp_len, const r1c_avg_job *jobs, int64_t n, int bit_depth) {
  if (!ctx || !dst) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, dst->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, dst->px_bytes);
  R1C_REQUIRE(ctx, n >= 0 && (n == 0 || (jobs && tmp1 && tmp2 && tmp_len > 0)), R1C_ERR_ARG, "null array");
  if (n == 0) return R1C_OK;
  for (int64_t i = 0; i < n; i++) {
    const r1c_avg_job &j = jobs[i];
    R1C_REQUIRE(ctx, j.w >= 1 && j.w <= 128 && j.h >= 1 && j.h <= 128, R1C_ERR_ARG, "avg job %lld: bad size", (long long)i);
    // asserts of mc_avg, src/mc.rs:621-632
    R1C_REQUIRE(ctx, j.tmp_off >= 0 && (int64_t)j.tmp_off + (int64_t)j.w * j.h <= tmp_len, R1C_ERR_BOUNDS, "avg job %lld: tmp is smaller than the compute block", (long long)i);
    R1C_REQUIRE(ctx, region_in_plane(dst->cfg, j.dst_x, j.dst_y, j.w, j.h), R1C_ERR_BOUNDS, "avg job %lld: dst is smaller than the compute block", (long long)i);
  }
  int rc;
  void *h_jobs, *d_jobs, *h_t, *d_t;
  const size_t jb = (size_t)n * sizeof(r1c_avg_job), tb = (size_t)tmp_len * 2;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, 2 * tb, &h_t))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, 2 * tb, &d_t))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_t, tmp1, tb);
  memcpy((uint8_t *)h_t + tb, tmp2, tb);
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_t, h_t, 2 * tb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_avg(ctx, dst, (const int16_t *)d_t, (const int16_t *)((uint8_t *)d_t + tb), (const r1c_avg_job *)d_jobs, n, bit_depth))) return rc;
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}
