// context.cu -- context, planes, host<->device staging, plane preparation
// kernels (pad / downsample).  Part of librav1e_cuda.so.
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

int r1c_set_err(r1c_ctx *ctx, int code, const char *fmt, ...) {
  if (ctx) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof ctx->err, fmt, ap);
    va_end(ap);
  }
  return code;
}

static const unsigned char TXW_[19] = {4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64};
static const unsigned char TXH_[19] = {4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16};

extern "C" int r1c_block_width(int b) { return (b >= 0 && b < R1C_BLOCK_SIZES_ALL) ? blk_w_of(b) : -1; }
extern "C" int r1c_block_height(int b) { return (b >= 0 && b < R1C_BLOCK_SIZES_ALL) ? blk_h_of(b) : -1; }
extern "C" int r1c_tx_width(int s) { return (s >= 0 && s < R1C_TX_SIZES_ALL) ? TXW_[s] : -1; }
extern "C" int r1c_tx_height(int s) { return (s >= 0 && s < R1C_TX_SIZES_ALL) ? TXH_[s] : -1; }

extern "C" r1c_ctx *r1c_create(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  r1c_ctx *ctx = (r1c_ctx *)calloc(1, sizeof(r1c_ctx));
  if (!ctx) return nullptr;
  ctx->device = device;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    free(ctx);
    return nullptr;
  }
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  return ctx;
}

extern "C" void r1c_destroy(r1c_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->peer) r1c_destroy(ctx->peer);
  if (ctx->win_ev_created) {
    cudaEventDestroy(ctx->win_ev[0]);
    cudaEventDestroy(ctx->win_ev[1]);
  }
  for (int i = 0; i < 4; i++)
    if (ctx->h_stage[i]) cudaFreeHost(ctx->h_stage[i]);
  for (int i = 0; i < 16; i++)
    if (ctx->d_scratch[i]) cudaFree(ctx->d_scratch[i]);
  for (int s = 0; s < R1C_PROF_SLOTS; s++)
    if (ctx->prof[s].created)
      for (int i = 0; i < R1C_PROF_RING; i++) {
        cudaEventDestroy(ctx->prof[s].ev[0][i]);
        cudaEventDestroy(ctx->prof[s].ev[1][i]);
      }
  if (ctx->frame_me.d_sbjobs) cudaFree(ctx->frame_me.d_sbjobs);
  if (ctx->frame_me.d_map) cudaFree(ctx->frame_me.d_map);
  if (ctx->frame_me.d_subjobs) cudaFree(ctx->frame_me.d_subjobs);
  cudaStreamDestroy(ctx->stream);
  free(ctx);
}

static bool host_pinned(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost) return true;
  (void)cudaGetLastError();
  return false;
}
int r1c_host_src(r1c_ctx *ctx, int slot, const void *p, size_t bytes, const void **src) {
  if (host_pinned(p)) {
    *src = p;
    return R1C_OK;
  }
  void *h;
  int rc = r1c_stage_host(ctx, slot, bytes, &h);
  if (rc) return rc;
  memcpy(h, p, bytes);
  *src = h;
  return R1C_OK;
}
int r1c_host_dst(r1c_ctx *ctx, int slot, void *p, size_t bytes, void **dst, int *staged) {
  if (host_pinned(p)) {
    *dst = p;
    *staged = 0;
    return R1C_OK;
  }
  *staged = 1;
  return r1c_stage_host(ctx, slot, bytes, dst);
}

extern "C" int r1c_sync(r1c_ctx *ctx) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}
extern "C" const char *r1c_last_error(const r1c_ctx *ctx) { return ctx ? ctx->err : "no context"; }
extern "C" void *r1c_stream(r1c_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
extern "C" int r1c_device(const r1c_ctx *ctx) { return ctx ? ctx->device : -1; }
extern "C" int64_t r1c_launch_count(const r1c_ctx *ctx) {
  return ctx ? ctx->launches + (ctx->peer ? ctx->peer->launches : 0) : 0;
}

// ---- per-kernel-class timing ---------------------------------------------------
static void prof_fold(r1c_ctx *ctx, int slot) {
  auto &p = ctx->prof[slot];
  if (p.n == 0) return;
  cudaEventSynchronize(p.ev[1][p.n - 1]);
  for (int i = 0; i < p.n; i++) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, p.ev[0][i], p.ev[1][i]) == cudaSuccess) p.ms_carry += ms;
  }
  p.n = 0;
}
void r1c_prof_begin(r1c_ctx *ctx, int slot) {
  if (!ctx->prof_on) return;
  auto &p = ctx->prof[slot];
  if (!p.created) {
    for (int i = 0; i < R1C_PROF_RING; i++) {
      cudaEventCreate(&p.ev[0][i]);
      cudaEventCreate(&p.ev[1][i]);
    }
    p.created = 1;
  }
  if (p.n == R1C_PROF_RING) prof_fold(ctx, slot);
  cudaEventRecord(p.ev[0][p.n], ctx->stream);
}
void r1c_prof_end(r1c_ctx *ctx, int slot) {
  if (!ctx->prof_on) return;
  auto &p = ctx->prof[slot];
  cudaEventRecord(p.ev[1][p.n], ctx->stream);
  p.n++;
  p.launches++;
}
extern "C" int r1c_profile(r1c_ctx *ctx, int enable) {
  if (!ctx) return R1C_ERR_ARG;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int s = 0; s < R1C_PROF_SLOTS; s++) {
    prof_fold(ctx, s);
    ctx->prof[s].ms_carry = 0;
    ctx->prof[s].launches = 0;
  }
  ctx->prof_on = enable != 0;
  return R1C_OK;
}
extern "C" int r1c_profile_read(r1c_ctx *ctx, int slot, double *ms, int64_t *launches) {
  if (!ctx || slot < 0 || slot >= R1C_PROF_SLOTS) return R1C_ERR_ARG;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  prof_fold(ctx, slot);
  if (ms) *ms = ctx->prof[slot].ms_carry;
  if (launches) *launches = ctx->prof[slot].launches;
  ctx->prof[slot].ms_carry = 0;
  ctx->prof[slot].launches = 0;
  return R1C_OK;
}

int r1c_stage_host(r1c_ctx *ctx, int slot, size_t bytes, void **out) {
  if (bytes > ctx->h_stage_cap[slot]) {
    if (ctx->h_stage[slot]) {
      R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      cudaFreeHost(ctx->h_stage[slot]);
      ctx->h_stage[slot] = nullptr;
      ctx->h_stage_cap[slot] = 0;
    }
    size_t cap = bytes + bytes / 4 + 4096;
    if (cudaMallocHost(&ctx->h_stage[slot], cap) != cudaSuccess)
      return r1c_set_err(ctx, R1C_ERR_NOMEM, "pinned staging of %zu bytes failed", cap);
    ctx->h_stage_cap[slot] = cap;
  }
  *out = ctx->h_stage[slot];
  return R1C_OK;
}

int r1c_scratch_dev(r1c_ctx *ctx, int slot, size_t bytes, void **out) {
  if (bytes > ctx->d_scratch_cap[slot]) {
    if (ctx->d_scratch[slot]) {
      R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      cudaFree(ctx->d_scratch[slot]);
      ctx->d_scratch[slot] = nullptr;
      ctx->d_scratch_cap[slot] = 0;
    }
    size_t cap = bytes + bytes / 4 + 4096;
    if (cudaMalloc(&ctx->d_scratch[slot], cap) != cudaSuccess)
      return r1c_set_err(ctx, R1C_ERR_NOMEM, "device scratch of %zu bytes failed", cap);
    ctx->d_scratch_cap[slot] = cap;
  }
  *out = ctx->d_scratch[slot];
  return R1C_OK;
}

// ---- planes ------------------------------------------------------------------
static int64_t align_pow2(int64_t v, int n) { return (v + (1ll << n) - 1) & ~((1ll << n) - 1); }

// Plane::new, src/frame/plane.rs:215-244 (STRIDE_ALIGNMENT_LOG2 = 5)
extern "C" int r1c_plane_cfg_new(r1c_plane_cfg *cfg, int64_t width, int64_t height, int64_t xdec,
                                 int64_t ydec, int64_t xpad, int64_t ypad, int px_bytes) {
  if (!cfg || width <= 0 || height <= 0 || xpad < 0 || ypad < 0 || (px_bytes != 1 && px_bytes != 2))
    return R1C_ERR_ARG;
  int al = 5 + 1 - px_bytes;
  cfg->xorigin = align_pow2(xpad, al);
  cfg->yorigin = ypad;
  cfg->stride = align_pow2(cfg->xorigin + width + xpad, al);
  cfg->alloc_height = cfg->yorigin + height + ypad;
  cfg->width = width;
  cfg->height = height;
  cfg->xdec = xdec;
  cfg->ydec = ydec;
  cfg->xpad = xpad;
  cfg->ypad = ypad;
  return R1C_OK;
}

static bool cfg_sane(const r1c_plane_cfg *c) {
  return c && c->stride > 0 && c->alloc_height > 0 && c->width > 0 && c->height > 0 &&
         c->xorigin >= 0 && c->yorigin >= 0 && c->xorigin + c->width <= c->stride &&
         c->yorigin + c->height <= c->alloc_height && c->stride < (1 << 20) &&
         c->alloc_height < (1 << 20);
}

template <typename T>
__global__ void fill_kernel(T *p, size_t n, T v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t step = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += step) p[i] = v;
}

extern "C" int r1c_plane_create(r1c_ctx *ctx, const r1c_plane_cfg *cfg, int px_bytes,
                                r1c_plane **out) {
  if (!ctx || !out) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, cfg_sane(cfg) && (px_bytes == 1 || px_bytes == 2), R1C_ERR_ARG, "bad plane config");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  r1c_plane *p = (r1c_plane *)calloc(1, sizeof(r1c_plane));
  if (!p) return R1C_ERR_NOMEM;
  p->cfg = *cfg;
  p->px_bytes = px_bytes;
  p->owned = true;
  p->device = ctx->device;
  size_t n = (size_t)cfg->stride * cfg->alloc_height;
  // +256 B tail so aligned vector loads that straddle the end stay inside the allocation
  if (cudaMalloc(&p->d_data, n * px_bytes + 256) != cudaSuccess) {
    free(p);
    return r1c_set_err(ctx, R1C_ERR_NOMEM, "cudaMalloc of plane (%zu bytes) failed", n * px_bytes);
  }
  int grid = ctx->sm_count * 8;
  if (px_bytes == 1) fill_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>((uint8_t *)p->d_data, n + 256, 128);
  else fill_kernel<uint16_t><<<grid, 256, 0, ctx->stream>>>((uint16_t *)p->d_data, n + 128, 128);
  R1C_LAUNCH_CHECK(ctx);
  *out = p;
  return R1C_OK;
}

extern "C" int r1c_plane_wrap(r1c_ctx *ctx, void *device_data, const r1c_plane_cfg *cfg,
                              int px_bytes, r1c_plane **out) {
  if (!ctx || !out || !device_data) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, cfg_sane(cfg) && (px_bytes == 1 || px_bytes == 2), R1C_ERR_ARG, "bad plane config");
  r1c_plane *p = (r1c_plane *)calloc(1, sizeof(r1c_plane));
  if (!p) return R1C_ERR_NOMEM;
  p->cfg = *cfg;
  p->px_bytes = px_bytes;
  p->owned = false;
  p->d_data = device_data;
  p->device = ctx->device;
  *out = p;
  return R1C_OK;
}

extern "C" void r1c_plane_destroy(r1c_plane *p) {
  if (!p) return;
  if (p->owned && p->d_data) {
    cudaSetDevice(p->device);
    cudaFree(p->d_data);
  }
  free(p);
}

extern "C" int r1c_plane_upload(r1c_ctx *ctx, r1c_plane *p, const void *host) {
  if (!ctx || !p || !host) return R1C_ERR_ARG;
  size_t bytes = (size_t)p->cfg.stride * p->cfg.alloc_height * p->px_bytes;
  R1C_CUDA(ctx, cudaMemcpyAsync(p->d_data, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_plane_upload_async(r1c_ctx *ctx, r1c_plane *p, const void *host) {
  if (!ctx || !p || !host) return R1C_ERR_ARG;
  size_t bytes = (size_t)p->cfg.stride * p->cfg.alloc_height * p->px_bytes;
  R1C_CUDA(ctx, cudaMemcpyAsync(p->d_data, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_plane_upload_rows_async(r1c_ctx *ctx, r1c_plane *p, const void *host, int64_t row_lo, int64_t row_hi) {
  if (!ctx || !p || !host) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, row_lo >= 0 && row_lo <= row_hi && row_hi <= p->cfg.alloc_height, R1C_ERR_ARG,
              "upload rows [%lld, %lld) outside the allocation of %lld rows", (long long)row_lo, (long long)row_hi,
              (long long)p->cfg.alloc_height);
  if (row_lo == row_hi) return R1C_OK;
  const size_t pitch = (size_t)p->cfg.stride * p->px_bytes;
  R1C_CUDA(ctx, cudaMemcpyAsync((uint8_t *)p->d_data + (size_t)row_lo * pitch, (const uint8_t *)host + (size_t)row_lo * pitch,
                                (size_t)(row_hi - row_lo) * pitch, cudaMemcpyHostToDevice, ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_plane_upload_rect_async(r1c_ctx *ctx, r1c_plane *p, const void *host, int64_t row_lo, int64_t row_hi,
                                           int64_t col_lo, int64_t col_hi) {
  if (!ctx || !p || !host) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, row_lo >= 0 && row_lo <= row_hi && row_hi <= p->cfg.alloc_height && col_lo >= 0 && col_lo <= col_hi && col_hi <= p->cfg.stride,
              R1C_ERR_ARG, "upload rect rows [%lld, %lld) x columns [%lld, %lld) outside the allocation of %lld x %lld", (long long)row_lo,
              (long long)row_hi, (long long)col_lo, (long long)col_hi, (long long)p->cfg.alloc_height, (long long)p->cfg.stride);
  if (row_lo == row_hi || col_lo == col_hi) return R1C_OK;
  const size_t pitch = (size_t)p->cfg.stride * p->px_bytes, off = (size_t)row_lo * pitch + (size_t)col_lo * p->px_bytes;
  R1C_CUDA(ctx, cudaMemcpy2DAsync((uint8_t *)p->d_data + off, pitch, (const uint8_t *)host + off, pitch, (size_t)(col_hi - col_lo) * p->px_bytes,
                                  (size_t)(row_hi - row_lo), cudaMemcpyHostToDevice, ctx->stream));
  return R1C_OK;
}

extern "C" int r1c_plane_download(r1c_ctx *ctx, const r1c_plane *p, void *host) {
  if (!ctx || !p || !host) return R1C_ERR_ARG;
  size_t bytes = (size_t)p->cfg.stride * p->cfg.alloc_height * p->px_bytes;
  R1C_CUDA(ctx, cudaMemcpyAsync(host, p->d_data, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return R1C_OK;
}

extern "C" void *r1c_plane_device_ptr(const r1c_plane *p) { return p ? p->d_data : nullptr; }
extern "C" const r1c_plane_cfg *r1c_plane_config(const r1c_plane *p) { return p ? &p->cfg : nullptr; }

// ---- Plane::pad, src/frame/plane.rs:269-314 ------------------------------------
// Two passes like the reference: (1) left/right replication on the `height`
// visible rows, (2) whole padded rows copied up / down.
template <typename T>
__global__ void pad_lr_kernel(T *data, int stride, int xorigin, int yorigin, int width, int height) {
  int y = blockIdx.x;
  if (y >= height) return;
  T *row = data + (size_t)(yorigin + y) * stride;
  T left = row[xorigin], right = row[xorigin + width - 1];
  for (int x = threadIdx.x; x < xorigin; x += blockDim.x) row[x] = left;
  for (int x = xorigin + width + threadIdx.x; x < stride; x += blockDim.x) row[x] = right;
}
template <typename T>
__global__ void pad_tb_kernel(T *data, int stride, int yorigin, int height, int alloc_h) {
  // grid.x enumerates the rows to fill: [0, yorigin) then [yorigin+height, alloc_h)
  int k = blockIdx.x;
  int y = k < yorigin ? k : (yorigin + height + (k - yorigin));
  if (y >= alloc_h) return;
  const T *src = data + (size_t)(k < yorigin ? yorigin : yorigin + height - 1) * stride;
  T *dst = data + (size_t)y * stride;
  for (int x = threadIdx.x; x < stride; x += blockDim.x) dst[x] = src[x];
}

extern "C" int r1c_plane_pad(r1c_ctx *ctx, r1c_plane *p, int64_t w, int64_t h) {
  if (!ctx || !p) return R1C_ERR_ARG;
  const r1c_plane_cfg &c = p->cfg;
  int width = (int)((w + c.xdec) >> c.xdec), height = (int)((h + c.ydec) >> c.ydec);
  R1C_REQUIRE(ctx, width > 0 && height > 0 && c.xorigin + width <= c.stride &&
                       c.yorigin + height <= c.alloc_height,
              R1C_ERR_BOUNDS, "pad(%lld,%lld) exceeds the allocation", (long long)w, (long long)h);
  int fill_rows = (int)(c.yorigin + (c.alloc_height - (c.yorigin + height)));
  if (p->px_bytes == 1) {
    pad_lr_kernel<uint8_t><<<height, 128, 0, ctx->stream>>>((uint8_t *)p->d_data, (int)c.stride, (int)c.xorigin, (int)c.yorigin, width, height);
    R1C_LAUNCH_CHECK(ctx);
    if (fill_rows > 0) {
      pad_tb_kernel<uint8_t><<<fill_rows, 256, 0, ctx->stream>>>((uint8_t *)p->d_data, (int)c.stride, (int)c.yorigin, height, (int)c.alloc_height);
      R1C_LAUNCH_CHECK(ctx);
    }
  } else {
    pad_lr_kernel<uint16_t><<<height, 128, 0, ctx->stream>>>((uint16_t *)p->d_data, (int)c.stride, (int)c.xorigin, (int)c.yorigin, width, height);
    R1C_LAUNCH_CHECK(ctx);
    if (fill_rows > 0) {
      pad_tb_kernel<uint16_t><<<fill_rows, 256, 0, ctx->stream>>>((uint16_t *)p->d_data, (int)c.stride, (int)c.yorigin, height, (int)c.alloc_height);
      R1C_LAUNCH_CHECK(ctx);
    }
  }
  return R1C_OK;
}

// ---- Plane::downsample_from, src/frame/plane.rs:399-423 --------------------------
template <typename T>
__global__ void downsample_kernel(T *dst, int dstride, int dxo, int dyo, int dw, int dh,
                                  const T *src, int sstride, int sxo, int syo) {
  int col = blockIdx.x * blockDim.x + threadIdx.x;
  int row = blockIdx.y;
  if (col >= dw || row >= dh) return;
  const T *s = src + (size_t)(syo + 2 * row) * sstride + sxo + 2 * col;
  uint32_t sum = (uint32_t)s[0] + s[1] + s[sstride] + s[sstride + 1];
  dst[(size_t)(dyo + row) * dstride + dxo + col] = (T)((sum + 2) >> 2);
}

extern "C" int r1c_plane_downsample(r1c_ctx *ctx, r1c_plane *dst, const r1c_plane *src) {
  if (!ctx || !dst || !src) return R1C_ERR_ARG;
  const r1c_plane_cfg &d = dst->cfg, &s = src->cfg;
  R1C_REQUIRE(ctx, d.width * 2 == s.width && d.height * 2 == s.height && dst->px_bytes == src->px_bytes,
              R1C_ERR_ARG, "downsample_from: source must be exactly 2x the destination");
  dim3 grid((unsigned)((d.width + 127) / 128), (unsigned)d.height);
  if (dst->px_bytes == 1)
    downsample_kernel<uint8_t><<<grid, 128, 0, ctx->stream>>>((uint8_t *)dst->d_data, (int)d.stride, (int)d.xorigin, (int)d.yorigin, (int)d.width, (int)d.height, (const uint8_t *)src->d_data, (int)s.stride, (int)s.xorigin, (int)s.yorigin);
  else
    downsample_kernel<uint16_t><<<grid, 128, 0, ctx->stream>>>((uint16_t *)dst->d_data, (int)d.stride, (int)d.xorigin, (int)d.yorigin, (int)d.width, (int)d.height, (const uint16_t *)src->d_data, (int)s.stride, (int)s.xorigin, (int)s.yorigin);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}
