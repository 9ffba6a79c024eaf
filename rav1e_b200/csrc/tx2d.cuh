// tx2d.cuh -- 2-D forward transform kernel (FwdTxfm2D::fht, reference
// src/transform/forward.rs:1804-1899) for one (W, H, column type, row type).
//
// Data movement per CTA: a slab of `BPC` consecutive blocks is contiguous in
// both the residual (i16) and the coefficient (i32) arrays, so it is moved
// with fully coalesced loads/stores and staged in shared memory:
//   global i16 slab --(128-bit loads)--> s_in
//   column pass : one thread per column, H-point transform in registers,
//                 s_in -> s_mid (row stride W+1 words: conflict-free for both
//                 the column-wise writes and the row-wise reads)
//   row pass    : one thread per row, W-point transform in registers,
//                 s_mid -> s_mid (in place, each thread owns its row)
//   s_mid --(coalesced 32-bit stores)--> global i32 slab
// Algorithmic traffic: 2 B in + 4 B out per coefficient (6 B/coeff), or
// 2*px_bytes in + 4 out when the residual is formed on the fly from two planes
// (`diff`, src/encoder.rs:1045-1059).
#pragma once

#include "common.cuh"
#include "tx1d.cuh"

namespace tx {

struct FhtArgs {
  const int16_t *residual;  // dense blocks (non-fused)
  int32_t *coeffs;
  long long n;              // number of blocks
  int sh0, sh1, sh2;        // FWD_SHIFT_* row for this bit depth (forward.rs:22-40)
  int col_flip, row_flip;
  // fused diff: residual = input - pred
  PlaneView input, pred;
  const r1c_tx_job *jobs;
};

// round_shift_array(arr, n, -entry), src/transform/mod.rs:499-521: entry > 0 is a
// wrapping left shift, entry < 0 a rounding right shift by -entry.
__device__ __forceinline__ int apply_shift(int v, int entry) {
  if (entry > 0) return (int)((unsigned)v << entry);
  if (entry < 0) return (int)((unsigned)v + (unsigned)((1 << -entry) >> 1)) >> -entry;
  return v;
}

template <int W, int H>
struct FhtCfg {
  static constexpr int MAXD = W > H ? W : H;
  // blocks per CTA: 256 threads for the small transforms, 128 for 32/64-point
  // ones (their register footprint is large)
  static constexpr int BPC = (MAXD >= 32 ? 128 : 256) / MAXD;
  static constexpr int THREADS = BPC * MAXD;
  static constexpr int MID_STRIDE = W + 1;
  static constexpr size_t SMEM = (size_t)BPC * W * H * 2 + (size_t)BPC * H * MID_STRIDE * 4;
};

template <int W, int H, int CT, int RT, int FUSED_PB>
__global__ void __launch_bounds__(FhtCfg<W, H>::THREADS) fht_kernel(FhtArgs a) {
  using C = FhtCfg<W, H>;
  constexpr int AREA = W * H;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int16_t *s_in = (int16_t *)smem_raw;
  int *s_mid = (int *)(smem_raw + (size_t)C::BPC * AREA * 2);

  const int tid = threadIdx.x;
  const long long blk0 = (long long)blockIdx.x * C::BPC;
  const int nb = (int)((a.n - blk0) < C::BPC ? (a.n - blk0) : C::BPC);

  // ---- stage the residual slab -------------------------------------------------
  if (FUSED_PB == 0) {
    const int16_t *src = a.residual + blk0 * AREA;
    const int total = nb * AREA;  // i16 elements, multiple of 16
    if ((((uintptr_t)src) & 15) == 0) {
      const int4 *s4 = (const int4 *)src;
      int4 *d4 = (int4 *)s_in;
      for (int i = tid; i < total / 8; i += C::THREADS) d4[i] = __ldg(s4 + i);
    } else {
      for (int i = tid; i < total; i += C::THREADS) s_in[i] = src[i];
    }
  } else {
    // diff(): residual[r*W + c] = input(in_x + c, in_y + r) - pred(pred_x + c, pred_y + r)
    for (int i = tid; i < nb * AREA; i += C::THREADS) {
      const int b = i / AREA, rc = i % AREA, r = rc / W, c = rc % W;
      const r1c_tx_job j = a.jobs[blk0 + b];
      int x, y;
      if (FUSED_PB == 1) {
        x = ((const uint8_t *)a.input.p0)[(long long)(j.in_y + r) * a.input.stride + j.in_x + c];
        y = ((const uint8_t *)a.pred.p0)[(long long)(j.pred_y + r) * a.pred.stride + j.pred_x + c];
      } else {
        x = ((const uint16_t *)a.input.p0)[(long long)(j.in_y + r) * a.input.stride + j.in_x + c];
        y = ((const uint16_t *)a.pred.p0)[(long long)(j.pred_y + r) * a.pred.stride + j.pred_x + c];
      }
      s_in[i] = (int16_t)(x - y);
    }
  }
  __syncthreads();

  // ---- columns (forward.rs:1826-1869) ----------------------------------------
  if (tid < nb * W) {
    const int b = tid / W, c = tid % W;
    const int16_t *col = s_in + b * AREA + c;
    int v[H], o[H];
#pragma unroll
    for (int r = 0; r < H; r++) {
      const int rr = a.col_flip ? (H - 1 - r) : r;  // :1829-1834 flip upside down
      v[r] = apply_shift((int)col[rr * W], a.sh0);
    }
    fwd1d<CT, H>(v, o);
    const int cc = a.row_flip ? (W - 1 - c) : c;  // intended semantic of :1858-1862
    int *dst = s_mid + b * H * C::MID_STRIDE + cc;
#pragma unroll
    for (int r = 0; r < H; r++) dst[r * C::MID_STRIDE] = apply_shift(o[r], a.sh1);
  }
  __syncthreads();

  // ---- rows (forward.rs:1872-1898) ---------------------------------------------
  if (tid < nb * H) {
    const int b = tid / H, r = tid % H;
    int *row = s_mid + (b * H + r) * C::MID_STRIDE;
    int v[W], o[W];
#pragma unroll
    for (int c = 0; c < W; c++) v[c] = row[c];
    fwd1d<RT, W>(v, o);
#pragma unroll
    for (int c = 0; c < W; c++) row[c] = apply_shift(o[c], a.sh2);
  }
  __syncthreads();

  // ---- coalesced store of the coefficient slab -------------------------------
  int32_t *dst = a.coeffs + blk0 * AREA;
  for (int i = tid; i < nb * AREA; i += C::THREADS) {
    const int rowi = i / W, c = i % W;  // rowi = b*H + r
    dst[i] = s_mid[rowi * C::MID_STRIDE + c];
  }
}

typedef int (*fht_launch_fn)(r1c_ctx *, const FhtArgs &, int fused_pb);

template <int W, int H, int CT, int RT>
int fht_launch(r1c_ctx *ctx, const FhtArgs &a, int fused_pb) {
  using C = FhtCfg<W, H>;
  if (a.n == 0) return R1C_OK;
  const unsigned grid = (unsigned)((a.n + C::BPC - 1) / C::BPC);
  if (C::SMEM > 48 * 1024) {
    static bool attr_set[3] = {false, false, false};
    if (!attr_set[fused_pb]) {
      cudaError_t e = cudaSuccess;
      if (fused_pb == 0) e = cudaFuncSetAttribute(fht_kernel<W, H, CT, RT, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
      else if (fused_pb == 1) e = cudaFuncSetAttribute(fht_kernel<W, H, CT, RT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
      else e = cudaFuncSetAttribute(fht_kernel<W, H, CT, RT, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
      if (e != cudaSuccess) return r1c_set_err(ctx, R1C_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      attr_set[fused_pb] = true;
    }
  }
  if (fused_pb == 0) fht_kernel<W, H, CT, RT, 0><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  else if (fused_pb == 1) fht_kernel<W, H, CT, RT, 1><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  else fht_kernel<W, H, CT, RT, 2><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// per-size table: [ct*3 + rt] -> launcher or nullptr (combination `unimplemented!`
// in the reference, forward.rs:1762-1768)
#define TX_ENTRY(W, H, CT, RT) \
  ((((CT) == TX1_ADST && (H) > 16) || ((RT) == TX1_ADST && (W) > 16) || ((CT) == TX1_ID && (H) > 32) || \
    ((RT) == TX1_ID && (W) > 32))                                                                       \
       ? (fht_launch_fn) nullptr                                                                        \
       : (fht_launch_fn)fht_launch<W, H, ((CT) == TX1_ADST && (H) > 16) ? TX1_DCT : ((CT) == TX1_ID && (H) > 32) ? TX1_DCT : (CT), \
                                   ((RT) == TX1_ADST && (W) > 16) ? TX1_DCT : ((RT) == TX1_ID && (W) > 32) ? TX1_DCT : (RT)>)

#define TX_SIZE_TABLE(NAME, W, H)                                                                      \
  extern const fht_launch_fn NAME[9];                                                                  \
  const fht_launch_fn NAME[9] = {TX_ENTRY(W, H, 0, 0), TX_ENTRY(W, H, 0, 1), TX_ENTRY(W, H, 0, 2),     \
                                 TX_ENTRY(W, H, 1, 0), TX_ENTRY(W, H, 1, 1), TX_ENTRY(W, H, 1, 2),     \
                                 TX_ENTRY(W, H, 2, 0), TX_ENTRY(W, H, 2, 1), TX_ENTRY(W, H, 2, 2)};

}  // namespace tx
