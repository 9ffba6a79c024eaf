// common.cuh -- internals shared by the translation units of librav1e_cuda.so.
// Product code: never includes anything from oracle/.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/rav1e_cuda.h"

struct r1c_plane {
  r1c_plane_cfg cfg;
  int px_bytes;
  void *d_data;   // whole allocation: stride * alloc_height pixels
  bool owned;
  int device;
};

#define R1C_PROF_RING 512
struct r1c_ctx {
  int device;
  cudaStream_t stream;
  char err[256];
  int64_t launches;
  int sm_count;
  // pinned host staging + device scratch, grown on demand (never shrinks)
  void *h_stage[4];
  size_t h_stage_cap[4];
  void *d_scratch[16];
  size_t d_scratch_cap[16];
  // per-kernel-class device timing (r1c_profile*): event pairs around the launches
  int prof_on;
  int prof_mc_slot;  // != 0: the MC launches belong to this slot (sub-pel path)
  int mc_taps_ready; // mc.cu: the packed filter table of this device has been written on this stream
  struct {
    cudaEvent_t ev[2][R1C_PROF_RING];
    int n;            // pairs recorded since the last read
    int created;
    double ms_carry;  // time of pairs already folded (ring wrapped)
    int64_t launches;
  } prof[R1C_PROF_SLOTS];
  // whole-frame motion estimation (me_frame.cu): job lists of the last
  // (frame size, parameters) seen, resident on the device
  struct {
    int valid;
    int32_t key[12];
    void *d_sbjobs, *d_map, *d_subjobs;
    int64_t n_sb, n_blocks;
  } frame_me;
  // r1c_frame_me_window*: a second context (stream + scratch of its own) that takes every other frame pair of a
  // window, created on first use, destroyed with this one; and the events that fork / join the two streams
  r1c_ctx *peer;
  cudaEvent_t win_ev[2];
  int win_ev_created;
};

// Device-side view of a plane: pointer to plane coordinate (0,0) + geometry.
struct PlaneView {
  const void *p0;   // address of pixel (0,0)
  int stride;       // pixels
  int xorigin, yorigin;
  int width, height;
  int alloc_w;      // = stride
  int alloc_h;
};

static inline PlaneView make_view(const r1c_plane *p) {
  PlaneView v;
  v.p0 = (const uint8_t *)p->d_data +
         ((size_t)p->cfg.yorigin * p->cfg.stride + p->cfg.xorigin) * p->px_bytes;
  v.stride = (int)p->cfg.stride;
  v.xorigin = (int)p->cfg.xorigin;
  v.yorigin = (int)p->cfg.yorigin;
  v.width = (int)p->cfg.width;
  v.height = (int)p->cfg.height;
  v.alloc_w = (int)p->cfg.stride;
  v.alloc_h = (int)p->cfg.alloc_height;
  return v;
}

// PlaneRegion::new bounds (src/tiling/plane_region.rs:139-142)
static inline bool region_in_plane(const r1c_plane_cfg &c, int64_t x, int64_t y, int64_t w,
                                   int64_t h) {
  return x >= -c.xorigin && y >= -c.yorigin && c.xorigin + x + w <= c.stride &&
         c.yorigin + y + h <= c.alloc_height;
}

int r1c_set_err(r1c_ctx *ctx, int code, const char *fmt, ...);
int r1c_stage_host(r1c_ctx *ctx, int slot, size_t bytes, void **out);
int r1c_scratch_dev(r1c_ctx *ctx, int slot, size_t bytes, void **out);
// Host side of a copy.  A page-locked caller buffer is the DMA source / target itself (*staged = 0); pageable memory
// goes through the pinned stage `slot`: r1c_host_src copies the caller's bytes in, r1c_host_dst hands out the stage and
// the caller copies out after synchronising (*staged = 1).
int r1c_host_src(r1c_ctx *ctx, int slot, const void *p, size_t bytes, const void **src);
int r1c_host_dst(r1c_ctx *ctx, int slot, void *p, size_t bytes, void **dst, int *staged);
// me_sb2.cu: superblock search for 8-bit and 10-bit planes; *handled = 0 -> caller falls back to me_sb.cu
int r1c_sb2_search(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_sb_params *p, int bit_depth,
                   const r1c_sb_job *d_jobs, int64_t n, r1c_me_result *d_results, const int2 *d_map, int64_t n_map,
                   int *handled);
// me_refine.cu: fused sub-pel tail of r1c_frame_me for 8-bit and 10-bit planes
int r1c_refine_frame(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_frame_me_params *p,
                     int bit_depth, const long long *off, const int *cols, const int *rows, const r1c_sb_job *d_sbs, int64_t n_sb,
                     r1c_me_result *d_results, int *handled);
// mc_tma.cu: put (kind 0) / prep (1) / dense put (2) through the TMA tile kernel; *handled = 0 -> caller uses mc.cu's kernel
int r1c_mc_tma_launch(r1c_ctx *ctx, int kind, r1c_plane *dst, int16_t *d_tmp, const r1c_plane *src, const r1c_mc_job *d_jobs,
                      int64_t n, int bit_depth, int *handled);
int r1c_mc_put_dense_dev(r1c_ctx *ctx, void *d_dense, const r1c_plane *src, const r1c_mc_job *d_jobs,
                         int64_t n, int bit_depth);

// tx.cu / itx.cu: per-size constants shared with enc_tx.cu (FWD_SHIFT_* row of forward.rs:22-40 for this bit depth;
// InvBlock::INTERMEDIATE_SHIFT, inverse.rs:1656-1665)
void r1c_fwd_shift_row(int tx_size, int bit_depth, int out[3]);
int r1c_inv_intermediate_shift(int tx_size);

// scoped event pair around a kernel launch (no-op unless r1c_profile(ctx, 1))
void r1c_prof_begin(r1c_ctx *ctx, int slot);
void r1c_prof_end(r1c_ctx *ctx, int slot);

#define R1C_CUDA(ctx, call)                                                              \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess)                                                               \
      return r1c_set_err((ctx), R1C_ERR_CUDA, "%s: %s (%s:%d)", #call,                   \
                         cudaGetErrorString(e_), __FILE__, __LINE__);                    \
  } while (0)

#define R1C_LAUNCH_CHECK(ctx)                                                            \
  do {                                                                                   \
    (ctx)->launches++;                                                                   \
    cudaError_t e_ = cudaGetLastError();                                                 \
    if (e_ != cudaSuccess)                                                               \
      return r1c_set_err((ctx), R1C_ERR_CUDA, "kernel launch: %s (%s:%d)",               \
                         cudaGetErrorString(e_), __FILE__, __LINE__);                    \
  } while (0)

#define R1C_REQUIRE(ctx, cond, code, ...)                                                \
  do {                                                                                   \
    if (!(cond)) return r1c_set_err((ctx), (code), __VA_ARGS__);                         \
  } while (0)

// BlockSize -> (w, h), order of src/partition.rs:116-140
// (log2(side) - 2 packed 4 bits per BlockSize: no table in local / constant memory)
__host__ __device__ inline int blk_w_of(int b) {
  const unsigned long long lo = 0x5544433322211100ull, hi = 0x423120ull;
  return 4 << (unsigned)((b < 16 ? lo >> (4 * b) : hi >> (4 * (b - 16))) & 15);
}
__host__ __device__ inline int blk_h_of(int b) {
  const unsigned long long lo = 0x5454343232121010ull, hi = 0x241302ull;
  return 4 << (unsigned)((b < 16 ? lo >> (4 * b) : hi >> (4 * (b - 16))) & 15);
}

static inline bool bit_depth_ok(int bd, int px_bytes) {
  return (bd == 8 && px_bytes == 1) || ((bd == 10 || bd == 12) && px_bytes == 2) ||
         (bd == 8 && px_bytes == 2);  // the reference's tests also run 8-bit data in u16 planes
}
