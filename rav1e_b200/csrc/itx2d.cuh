// itx2d.cuh -- 2-D inverse transform + add (NativeInvTxfm2D::inv_txfm2d / inv_txfm2d_add, reference
// src/transform/inverse.rs:1939-2107) for one (W, H, column type, row type).
//
// A CTA of 128 threads takes `BPC` = 128 / W consecutive blocks of the call:
//   row pass    : one thread per coded row (the first min(H,32) rows; the first min(W,32) coefficients of
//                 each, the rest are zero: 64-point transforms signal 32), 16-byte loads straight from the
//                 coefficient array (a thread owns its row's contiguous bytes), the rectangular 1/sqrt2
//                 scale and the bd+8 clamp, the W-point transform in registers, then the intermediate
//                 rounding shift and the max(bd+6,16) clamp on the way into shared memory
//                 (row stride W+1 words: conflict-free for these row-wise writes and the column reads)
//   column pass : one thread per column, H-point transform in registers, round_shift(.., 4), and the
//                 read-modify-write of the destination pixels (threads of a block touch consecutive pixels)
// Algorithmic traffic per coefficient: 4 B of coefficients (coded ones) + px_bytes read + px_bytes written.
#pragma once

#include <type_traits>

#include "common.cuh"
#include "itx1d.cuh"

namespace itx {

struct ItxArgs {
  const int32_t *coeffs;  // block i at coeffs + i * coded
  PlaneView dst;
  const r1c_itx_job *jobs;
  long long n;
  Range rg_row, rg_col;  // clamp_value bounds for bd + 8 and max(bd + 6, 16) bits
  int sh, rnd;           // InvBlock::INTERMEDIATE_SHIFT and its rounding term
  int maxv;              // (1 << bd) - 1
  int col_flip, row_flip;
  int in16;              // coefficient array is 16-byte aligned
  // fused dequantize (src/quantize.rs:318-334): the array holds qcoeffs, each one is scaled on load
  int dq_on, dq_lts;
  unsigned dq_dc, dq_ac;
};

template <int W, int H>
struct ItxCfg {
  static constexpr int WI = W < 32 ? W : 32, HI = H < 32 ? H : 32;
  static constexpr int CODED = WI * HI;
  static constexpr int MAXD = W > H ? W : H;
  // blocks per CTA: as many as give every thread exactly one column (the H-point transform + the pixel
  // read-modify-write; the coded rows are looped over or leave threads idle), except for the 4:1 wide blocks
  // (16x4, 32x8, 64x16), where 3 of 4 warps would idle through the long row transforms: those are sized by the
  // coded rows and loop over the columns (measured: 2:1 blocks and the 64-wide ones with 32 coded rows lose)
  static constexpr int THREADS = 128;
  static constexpr int BPC = THREADS / (W >= 4 * HI ? HI : W);  // blocks per CTA
  static constexpr bool ROWS_ONCE = BPC * HI <= THREADS;
  static constexpr bool COLS_ONCE = BPC * W <= THREADS;
  static constexpr int STRIDE = W + 1;
  static constexpr int BLK = HI * STRIDE;  // words per block: the coded rows only
  static constexpr size_t SMEM = (size_t)BPC * BLK * 4;
};

template <int W, int H, int CT, int RT, int PB>
__global__ void __launch_bounds__(ItxCfg<W, H>::THREADS) itx_kernel(const __grid_constant__ ItxArgs a) {
  using C = ItxCfg<W, H>;
  constexpr bool RECT = log2c(W) - log2c(H) == 1 || log2c(H) - log2c(W) == 1;  // get_rect_tx_log_ratio, :2070
  extern __shared__ __align__(16) int s_mid[];
  const int tid = threadIdx.x;
  const int nslabs = (int)((a.n + C::BPC - 1) / C::BPC);
  for (int slab = blockIdx.x; slab < nslabs; slab += gridDim.x) {
    const long long blk0 = (long long)slab * C::BPC;
    const int nb = (int)min((long long)C::BPC, a.n - blk0);

    using Px = typename std::conditional<PB == 1, uint8_t, uint16_t>::type;
    // the destination pixels are read in batches of up to 32 independent loads (a load placed after a store
    // through another pointer could not be moved above it by the compiler: one DRAM latency per row otherwise);
    // blocks up to 32 rows issue them here, so that they are in flight during both transforms
    constexpr int CH = H < 32 ? H : 32;
    constexpr bool PREFETCH = C::COLS_ONCE && H <= 32;  // one column per thread: its pixels are fetched before the row pass
    const bool col_live = tid < nb * W;
    Px *p = nullptr;
    int px[CH];
    if (C::COLS_ONCE && col_live) {
      const r1c_itx_job j = a.jobs[blk0 + tid / W];
      p = (Px *)a.dst.p0 + ((long long)j.y * a.dst.stride + j.x + tid % W);
      if (PREFETCH) {
#pragma unroll
        for (int r = 0; r < H; r++) px[r] = p[(long long)r * a.dst.stride];
      }
    }

    // ---- rows (:2073-2092) -----------------------------------------------------------
    auto row_item = [&](int it) {
      const int b = it / C::HI, r = it % C::HI;
      const int32_t *src = a.coeffs + (blk0 + b) * C::CODED + r * C::WI;
      int v[W], o[W];
      if (a.in16) {
#pragma unroll
        for (int c = 0; c < C::WI; c += 4) {
          const int4 q = __ldg((const int4 *)(src + c));
          v[c] = q.x, v[c + 1] = q.y, v[c + 2] = q.z, v[c + 3] = q.w;
        }
      } else {
#pragma unroll
        for (int c = 0; c < C::WI; c++) v[c] = __ldg(src + c);
      }
      if (a.dq_on) {  // *r = (c * quant + ((c >> 31) & offset)) >> log_tx_scale, wrapping i32 product
        const unsigned off = (1u << a.dq_lts) - 1u;
#pragma unroll
        for (int c = 0; c < C::WI; c++)
          v[c] = (int)((unsigned)v[c] * ((c == 0 && r == 0) ? a.dq_dc : a.dq_ac) + ((unsigned)(v[c] >> 31) & off)) >> a.dq_lts;
      }
      int l1 = 0;  // sum |input|: decides whether the transform's clamps can change anything (itx1d.cuh)
#pragma unroll
      for (int c = 0; c < W; c++) {
        if (c >= C::WI) v[c] = 0;
        else {
          if (RECT) v[c] = (int)((unsigned)v[c] * 2896u + 2048u) >> 12;  // round_shift(raw * INV_SQRT2, SQRT2_BITS)
          l1 = (int)min((unsigned)l1 + (unsigned)abs(v[c]), 0x40000000u);  // saturating: |INT_MIN| and overflow stay large
        }
      }
      // the input clamp and the clamps inside are no-ops when the row's L1 norm stays inside the range
      if (__all_sync(__activemask(), l1 + NOCLAMP_MARGIN <= a.rg_row.hi)) {
        inv1d<RT, W, false>(v, o, a.rg_row);
      } else {
#pragma unroll
        for (int c = 0; c < C::WI; c++) v[c] = cl(v[c], a.rg_row);
        inv1d<RT, W, true>(v, o, a.rg_row);
      }
      int *dst = s_mid + b * C::BLK + r * C::STRIDE;
      if (a.row_flip) {  // av1_iflipadst*: the reversed output of the ADST
#pragma unroll
        for (int c = 0; c < W; c++) dst[W - 1 - c] = cl((o[c] + a.rnd) >> a.sh, a.rg_col);
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) dst[c] = cl((o[c] + a.rnd) >> a.sh, a.rg_col);
      }
    };
    if constexpr (C::ROWS_ONCE) {
      if (tid < nb * C::HI) row_item(tid);
    } else {
      for (int it = tid; it < nb * C::HI; it += C::THREADS) row_item(it);
    }
    __syncthreads();

    // ---- columns (:2095-2106) and the add (:2017-2028) ------------------------------------
    auto col_item = [&](int it) {
      const int b = it / W, c = it % W;
      if (!C::COLS_ONCE) {
        const r1c_itx_job j = a.jobs[blk0 + b];
        p = (Px *)a.dst.p0 + ((long long)j.y * a.dst.stride + j.x + c);
      }
      const int *col = s_mid + b * C::BLK + c;
      int v[H], o[H];
      int l1 = 0;
#pragma unroll
      for (int r = 0; r < H; r++) {
        v[r] = r < C::HI ? col[r * C::STRIDE] : 0;
        if (r < C::HI) l1 += abs(v[r]);  // |v| <= 2^15 after the row pass' clamp: no overflow
      }
      if (__all_sync(__activemask(), l1 + NOCLAMP_MARGIN <= a.rg_col.hi)) inv1d<CT, H, false>(v, o, a.rg_col);
      else inv1d<CT, H, true>(v, o, a.rg_col);
#pragma unroll
      for (int r0 = 0; r0 < H; r0 += CH) {
        if (!PREFETCH) {
#pragma unroll
          for (int r = 0; r < CH; r++) px[r] = p[(long long)(r0 + r) * a.dst.stride];
        }
#pragma unroll
        for (int r = 0; r < CH; r++) {
          // av1_iflipadst*: output row rr of the block comes from o[H - 1 - rr]
          const int t = ((a.col_flip ? o[H - 1 - (r0 + r)] : o[r0 + r]) + 8) >> 4;
          p[(long long)(r0 + r) * a.dst.stride] = (Px)min(max(px[r] + t, 0), a.maxv);
        }
      }
    };
    if constexpr (C::COLS_ONCE) {
      if (col_live) col_item(tid);
    } else {
      for (int it = tid; it < nb * W; it += C::THREADS) col_item(it);
    }
    __syncthreads();  // s_mid is rewritten by the next slab's row pass
  }
}

typedef int (*itx_launch_fn)(r1c_ctx *, const ItxArgs &, int px_bytes);

template <int W, int H, int CT, int RT, int PB>
int itx_launch_one(r1c_ctx *ctx, ItxArgs a) {
  using C = ItxCfg<W, H>;
  static int per_sm = 0;  // resident CTAs per SM of this instantiation
  if (!per_sm) {
    int nb = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, itx_kernel<W, H, CT, RT, PB>, C::THREADS, C::SMEM);
    if (e != cudaSuccess || nb < 1) return r1c_set_err(ctx, R1C_ERR_CUDA, "itx kernel does not fit an SM: %s", cudaGetErrorString(e));
    per_sm = nb;
  }
  const long long nslabs = (a.n + C::BPC - 1) / C::BPC;
  const long long full = (long long)ctx->sm_count * per_sm;
  const unsigned grid = (unsigned)(nslabs < full ? nslabs : full);
  a.in16 = (((uintptr_t)a.coeffs) & 15) == 0;
  itx_kernel<W, H, CT, RT, PB><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

template <int W, int H, int CT, int RT>
int itx_launch(r1c_ctx *ctx, const ItxArgs &a, int px_bytes) {
  if (a.n == 0) return R1C_OK;
  if (px_bytes == 1) return itx_launch_one<W, H, CT, RT, 1>(ctx, a);
  return itx_launch_one<W, H, CT, RT, 2>(ctx, a);
}

// per-size table: [ct*3 + rt] -> launcher, or nullptr where the reference has no 1-D function
// (INV_TXFM_FNS, inverse.rs:1610-1634: ADST up to 16 points, identity up to 32)
#define ITX_BAD(T, N) (((T) == I1_ADST && (N) > 16) || ((T) == I1_ID && (N) > 32))
#define ITX_ENTRY(W, H, CT, RT)                                  \
  ((ITX_BAD(CT, H) || ITX_BAD(RT, W)) ? (itx_launch_fn) nullptr \
                                      : (itx_launch_fn)itx_launch<W, H, ITX_BAD(CT, H) ? I1_DCT : (CT), ITX_BAD(RT, W) ? I1_DCT : (RT)>)

#define ITX_SIZE_TABLE(NAME, W, H)                                                                       \
  extern const itx_launch_fn NAME[9];                                                                    \
  const itx_launch_fn NAME[9] = {ITX_ENTRY(W, H, 0, 0), ITX_ENTRY(W, H, 0, 1), ITX_ENTRY(W, H, 0, 2),    \
                                 ITX_ENTRY(W, H, 1, 0), ITX_ENTRY(W, H, 1, 1), ITX_ENTRY(W, H, 1, 2),    \
                                 ITX_ENTRY(W, H, 2, 0), ITX_ENTRY(W, H, 2, 1), ITX_ENTRY(W, H, 2, 2)};

}  // namespace itx
