// enc_tx.cuh -- the residual chain of encode_tx_block (reference src/encoder.rs:1143-1224) for one
// (W, H, column type, row type), W, H <= 32, as ONE kernel:
//   diff (:1045-1059) -> forward_transform (src/transform/forward.rs:1804-1899) -> QuantizationContext::quantize
//   (src/quantize.rs:257-316) -> dequantize (:318-334) -> inverse_transform_add (src/transform/inverse.rs:1939-2107)
//   -> tx-domain distortion (src/encoder.rs:1210-1224)
// The coefficients of a block never leave the SM: a CTA takes BPC consecutive jobs, their W x H i32 coefficients
// live in one shared-memory array (rows W + 1 words apart: conflict-free for the column-wise and the row-wise
// accesses of all four 1-D passes) from the first column pass to the last one.  Global traffic per pixel:
//   px_bytes (input) + px_bytes (prediction, read from the reconstruction plane) + 4 (qcoeffs, written for the
//   entropy coder) + px_bytes (reconstruction written back) [+ 8 / (W H) tx_dist]
// against 2 px + 4 | 4 + 4 (+4) | 4 + 4 | 4 + 2 px for the four separate kernels.
//   phase 1  forward columns : one thread per column; pixels of both planes straight from global memory
//   phase 2  forward rows    : one thread per row, in place
//            (BPC = THREADS / min(W, H): the pass with the long 1-D transforms has exactly one item per thread, the
//            pass with the short ones max / min times as many and loops)
//   phase 3  quantize        : LPB lanes per block walk the scan (quant_dev.cuh), in place; tx_dist
//   phase 4  qcoeffs out (coalesced), then inverse rows with the dequantizer on the load, in place
//   phase 5  inverse columns + add onto the prediction + clamp, pixels stored
#pragma once

#include "common.cuh"
#include "itx1d.cuh"
#include "quant_dev.cuh"
#include "tx1d.cuh"

namespace enc {

struct EncArgs {
  PlaneView input, rec;
  const r1c_tx_job *jobs;  // (in_x, in_y) in `input`, (pred_x, pred_y) in `rec`
  long long n;
  int32_t *qcoeffs;   // block i at qcoeffs + i * W * H
  uint64_t *tx_dist;  // or nullptr
  int reconstruct;    // run phases 4b / 5 (`!fi.use_tx_domain_distortion || need_recon_pixel`, src/encoder.rs:1203)
  int px_bytes;
  int ls0, rs1, rnd1, ls2, rs2, rnd2;  // FWD_SHIFT_* of (tx_size, bit depth) as branch-free amounts, see tx2d.cuh
  int col_flip, row_flip;
  quant::QuantConsts q;
  int dist_bits;
  itx::Range rg_row, rg_col;  // clamp_value bounds for bd + 8 and max(bd + 6, 16) bits
  int ish, irnd;              // InvBlock::INTERMEDIATE_SHIFT and its rounding term
  int maxv;                   // (1 << bd) - 1
};

template <int W, int H>
struct EncCfg {
  static_assert(W <= 32 && H <= 32, "64-point transforms go through the separate kernels");
  static constexpr int MAXD = W > H ? W : H, MIND = W < H ? W : H;
  static constexpr int THREADS = MAXD >= 32 ? 128 : 256;
  // blocks per CTA: the pass with the LONG transforms (columns of a tall block, rows of a wide one) has THREADS items,
  // one per thread; the other pass has MAXD / MIND times as many short ones and loops
  static constexpr int BPC = THREADS / MIND;
  static constexpr int STRIDE = W + 1;
  static constexpr int BLK = H * STRIDE;  // words per block in shared memory
  static constexpr int CODED = W * H;
  // lanes per block / chunks per lane of the quantizer: the table of quant.cu
  static constexpr int LPB = CODED >= 128 ? 32 : (CODED == 16 ? 4 : 8);
  static constexpr int NCH = CODED / LPB;
  static constexpr int BPW = 32 / LPB;
  static constexpr int GROUPS = BPC / BPW;
  static_assert(BPC % BPW == 0, "a warp's blocks must belong to one CTA");
  static constexpr int LOGW = W == 4 ? 2 : W == 8 ? 3 : W == 16 ? 4 : 5;
};

template <typename Px>
__device__ __forceinline__ int ld_px(const void *base, long long idx) {
  return (int)((const Px *)base)[idx];
}

template <int W, int H, int CT, int RT>
__global__ void __launch_bounds__(EncCfg<W, H>::THREADS) enc_tx_kernel(const __grid_constant__ EncArgs a) {
  using C = EncCfg<W, H>;
  constexpr bool RECT = C::LOGW - (H == 4 ? 2 : H == 8 ? 3 : H == 16 ? 4 : 5) == 1 ||
                        (H == 4 ? 2 : H == 8 ? 3 : H == 16 ? 4 : 5) - C::LOGW == 1;  // get_rect_tx_log_ratio, inverse.rs:2070
  __shared__ int s_c[C::BPC * C::BLK];
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  const int nslabs = (int)((a.n + C::BPC - 1) / C::BPC);
  const int lts = (int)a.q.qc.log_tx_scale;
  for (int slab = blockIdx.x; slab < nslabs; slab += gridDim.x) {
    const long long blk0 = (long long)slab * C::BPC;
    const int nb = (int)min((long long)C::BPC, a.n - blk0);

    // ---- 1: diff + forward columns (forward.rs:1826-1869) ---------------------------------------
#pragma unroll 1
    for (int it = tid; it < nb * W; it += C::THREADS) {
      const int b = it / W, c = it % W;
      const r1c_tx_job j = a.jobs[blk0 + b];
      const long long in_off = (long long)j.in_y * a.input.stride + j.in_x + c;
      const long long rec_off = (long long)j.pred_y * a.rec.stride + j.pred_x + c;
      int v[H], o[H];
      if (a.px_bytes == 1) {
#pragma unroll
        for (int r = 0; r < H; r++)
          v[r] = ld_px<uint8_t>(a.input.p0, in_off + (long long)r * a.input.stride) - ld_px<uint8_t>(a.rec.p0, rec_off + (long long)r * a.rec.stride);
      } else {
#pragma unroll
        for (int r = 0; r < H; r++)
          v[r] = ld_px<uint16_t>(a.input.p0, in_off + (long long)r * a.input.stride) - ld_px<uint16_t>(a.rec.p0, rec_off + (long long)r * a.rec.stride);
      }
      if (a.col_flip) {  // :1829-1834 flip upside down
#pragma unroll
        for (int r = 0; r < H / 2; r++) {
          const int t = v[r];
          v[r] = v[H - 1 - r], v[H - 1 - r] = t;
        }
      }
#pragma unroll
      for (int r = 0; r < H; r++) v[r] = (int)((unsigned)(int)(int16_t)v[r] << a.ls0);
      tx::fwd1d<CT, H>(v, o);
      const int cc = a.row_flip ? (W - 1 - c) : c;
      int *dst = s_c + b * C::BLK + cc;
      if (a.rs1) {
#pragma unroll
        for (int r = 0; r < H; r++) dst[r * C::STRIDE] = (int)((unsigned)o[r] + (unsigned)a.rnd1) >> a.rs1;
      } else {
#pragma unroll
        for (int r = 0; r < H; r++) dst[r * C::STRIDE] = o[r];
      }
    }
    __syncthreads();

    // ---- 2: forward rows (forward.rs:1872-1898), in place ----------------------------------------
#pragma unroll 1
    for (int it = tid; it < nb * H; it += C::THREADS) {  // it = b * H + r: rows of the slab are STRIDE words apart
      int *row = s_c + it * C::STRIDE;
      int v[W], o[W];
#pragma unroll
      for (int c = 0; c < W; c++) v[c] = row[c];
      tx::fwd1d<RT, W>(v, o);
      if (a.rs2) {
#pragma unroll
        for (int c = 0; c < W; c++) row[c] = (int)((unsigned)o[c] + (unsigned)a.rnd2) >> a.rs2;
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) row[c] = (int)((unsigned)o[c] << a.ls2);
      }
    }
    __syncthreads();

    // ---- 3: quantize in place (quantize.rs:257-316) + tx-domain distortion (encoder.rs:1210-1224) --
    for (int g = wrp; g < C::GROUPS && g * C::BPW < nb; g += C::THREADS / 32) {
      const int b = g * C::BPW + lane / C::LPB;  // blocks past nb (same warp) go through the motions on stale words
      unsigned long long dist = quant::quantize_lanes<C::LPB, C::NCH, C::LOGW>(s_c + b * C::BLK, lane & (C::LPB - 1),
                                                                               lane & ~(C::LPB - 1), a.q, a.tx_dist != nullptr);
      if (a.tx_dist) {
#pragma unroll
        for (int m = C::LPB >> 1; m; m >>= 1) dist += __shfl_xor_sync(0xffffffffu, dist, m, C::LPB);
        if ((lane & (C::LPB - 1)) == 0 && b < nb) a.tx_dist[blk0 + b] = (dist + (1ull << (a.dist_bits - 1))) >> a.dist_bits;
      }
    }
    __syncthreads();

    // ---- 4a: quantized coefficients out, coalesced ------------------------------------------------
    {
      int32_t *out = a.qcoeffs + blk0 * C::CODED;
      if (nb == C::BPC) {
        constexpr int ITER = C::BPC * C::CODED / C::THREADS;
        constexpr int ROWS_PER_ITER = C::THREADS / W;
        // row index = b * H + r over the whole slab; block b's rows start at b * BLK = b * H * STRIDE: uniform pitch
        const int *src = s_c + (tid / W) * C::STRIDE + (tid % W);
#pragma unroll
        for (int it = 0; it < ITER; it++) out[it * C::THREADS + tid] = src[it * ROWS_PER_ITER * C::STRIDE];
      } else {
        for (int i = tid; i < nb * C::CODED; i += C::THREADS) out[i] = s_c[(i / W) * C::STRIDE + (i % W)];
      }
    }
    if (!a.reconstruct) {
      __syncthreads();  // s_c is rewritten by the next slab's column pass
      continue;
    }
    // a row is read here (above, by any thread) and rewritten below by its own thread only
    __syncthreads();

    // ---- 4b: dequantize (quantize.rs:318-334) + inverse rows (inverse.rs:2073-2092), in place -----
#pragma unroll 1
    for (int it = tid; it < nb * H; it += C::THREADS) {
      int *row = s_c + it * C::STRIDE;
      const int r = it % H;
      int v[W], o[W], l1 = 0;
      const unsigned off = (1u << lts) - 1u;
#pragma unroll
      for (int c = 0; c < W; c++) {
        const int qv = row[c];
        int t = (int)((unsigned)qv * ((c == 0 && r == 0) ? a.q.qc.dc_quant : a.q.qc.ac_quant) + ((unsigned)(qv >> 31) & off)) >> lts;
        if (RECT) t = (int)((unsigned)t * 2896u + 2048u) >> 12;  // round_shift(raw * INV_SQRT2, SQRT2_BITS)
        v[c] = t;
        l1 = (int)min((unsigned)l1 + (unsigned)abs(t), 0x40000000u);
      }
      // the input clamp and the clamps inside the transform are no-ops when the row's L1 norm stays inside the range
      // (itx1d.cuh, NOCLAMP_MARGIN): the usual case for the residual of a prediction
      if (__all_sync(__activemask(), l1 + itx::NOCLAMP_MARGIN <= a.rg_row.hi)) {
        itx::inv1d<RT, W, false>(v, o, a.rg_row);
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) v[c] = itx::cl(v[c], a.rg_row);
        itx::inv1d<RT, W, true>(v, o, a.rg_row);
      }
      if (a.row_flip) {
#pragma unroll
        for (int c = 0; c < W; c++) row[W - 1 - c] = itx::cl((o[c] + a.irnd) >> a.ish, a.rg_col);
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) row[c] = itx::cl((o[c] + a.irnd) >> a.ish, a.rg_col);
      }
    }
    __syncthreads();

    // ---- 5: inverse columns (inverse.rs:2095-2106) and the add (:2017-2028) -------------------------
#pragma unroll 1
    for (int it = tid; it < nb * W; it += C::THREADS) {
      const int b = it / W, c = it % W;
      const r1c_tx_job j = a.jobs[blk0 + b];
      const long long rec_off = (long long)j.pred_y * a.rec.stride + j.pred_x + c;
      const int *col = s_c + b * C::BLK + c;
      int v[H], o[H];
      int l1 = 0;
#pragma unroll
      for (int r = 0; r < H; r++) {
        v[r] = col[r * C::STRIDE];
        l1 += abs(v[r]);  // |v| < 2^17 after the row pass' clamp
      }
      if (__all_sync(__activemask(), l1 + itx::NOCLAMP_MARGIN <= a.rg_col.hi)) itx::inv1d<CT, H, false>(v, o, a.rg_col);
      else itx::inv1d<CT, H, true>(v, o, a.rg_col);
      if (a.px_bytes == 1) {
        uint8_t *p = (uint8_t *)a.rec.p0 + rec_off;
        int px[H];
#pragma unroll
        for (int r = 0; r < H; r++) px[r] = p[(long long)r * a.rec.stride];
#pragma unroll
        for (int r = 0; r < H; r++) {
          const int t = ((a.col_flip ? o[H - 1 - r] : o[r]) + 8) >> 4;
          p[(long long)r * a.rec.stride] = (uint8_t)min(max(px[r] + t, 0), a.maxv);
        }
      } else {
        uint16_t *p = (uint16_t *)a.rec.p0 + rec_off;
        int px[H];
#pragma unroll
        for (int r = 0; r < H; r++) px[r] = p[(long long)r * a.rec.stride];
#pragma unroll
        for (int r = 0; r < H; r++) {
          const int t = ((a.col_flip ? o[H - 1 - r] : o[r]) + 8) >> 4;
          p[(long long)r * a.rec.stride] = (uint16_t)min(max(px[r] + t, 0), a.maxv);
        }
      }
    }
    __syncthreads();  // s_c is rewritten by the next slab's column pass
  }
}

typedef int (*enc_launch_fn)(r1c_ctx *, const EncArgs &);

template <int W, int H, int CT, int RT>
int enc_launch(r1c_ctx *ctx, const EncArgs &a) {
  using C = EncCfg<W, H>;
  if (a.n == 0) return R1C_OK;
  static int per_sm = 0;  // resident CTAs per SM of this instantiation
  if (!per_sm) {
    int nb = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, enc_tx_kernel<W, H, CT, RT>, C::THREADS, 0);
    if (e != cudaSuccess || nb < 1) return r1c_set_err(ctx, R1C_ERR_CUDA, "enc_tx kernel does not fit an SM: %s", cudaGetErrorString(e));
    per_sm = nb;
  }
  const long long nslabs = (a.n + C::BPC - 1) / C::BPC;
  const long long full = (long long)ctx->sm_count * per_sm;
  const unsigned grid = (unsigned)(nslabs < full ? nslabs : full);
  enc_tx_kernel<W, H, CT, RT><<<grid, C::THREADS, 0, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// per-size table: [ct*3 + rt] -> launcher, or nullptr where the reference has no 1-D function (forward.rs:1762-1768,
// inverse.rs:1610-1634: ADST up to 16 points; identity up to 32)
#define ENC_BAD(T, N) ((T) == 1 && (N) > 16)
#define ENC_ENTRY(W, H, CT, RT)                                   \
  ((ENC_BAD(CT, H) || ENC_BAD(RT, W)) ? (enc_launch_fn) nullptr  \
                                      : (enc_launch_fn)enc_launch<W, H, ENC_BAD(CT, H) ? 0 : (CT), ENC_BAD(RT, W) ? 0 : (RT)>)

#define ENC_SIZE_TABLE(NAME, W, H)                                                                      \
  extern const enc_launch_fn NAME[9];                                                                   \
  const enc_launch_fn NAME[9] = {ENC_ENTRY(W, H, 0, 0), ENC_ENTRY(W, H, 0, 1), ENC_ENTRY(W, H, 0, 2),   \
                                 ENC_ENTRY(W, H, 1, 0), ENC_ENTRY(W, H, 1, 1), ENC_ENTRY(W, H, 1, 2),   \
                                 ENC_ENTRY(W, H, 2, 0), ENC_ENTRY(W, H, 2, 1), ENC_ENTRY(W, H, 2, 2)};

}  // namespace enc
