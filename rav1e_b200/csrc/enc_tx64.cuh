// enc_tx64.cuh -- the residual chain of encode_tx_block (reference src/encoder.rs:1143-1224) as ONE kernel for the
// five transform sizes with a 64-point side (TX_64X64, 32X64, 64X32, 16X64, 64X16; the 64-point side is a DCT, the
// other side any 1-D transform the reference has for its length).  Same phases as enc_tx.cuh; what differs is the
// TWO layouts of this fork:
//   * `fht` leaves dense rows of W coefficients (src/transform/forward.rs:1885-1889, no 64 -> 32 repack);
//   * the quantizer, the dequantizer, tx_dist and the inverse transform see the first CODED = min(W,32) * min(H,32)
//     words of that buffer, the inverse as rows of min(W,32) (`input.chunks(width.min(32))`,
//     src/transform/inverse.rs:2073-2092).
// So only the first FR = CODED / W dense rows of the forward transform are ever read: the column pass keeps those
// outputs only (the compiler drops the rest of the butterflies) and the row pass runs on FR rows.  Shared memory holds
// max(FR, CH) rows of W + 1 words per block: the dense forward rows first, then -- after a barrier between the loads
// and the stores of the inverse row pass, whose output row r overwrites the inputs of rows 2r and 2r + 1 -- the
// CH = min(H,32) intermediate rows of the inverse.
#pragma once

#include "enc_tx.cuh"

#ifndef ENC64_GROUPS
#define ENC64_GROUPS 2
#endif

namespace enc {

template <int W, int H>
struct Enc64Cfg {
  static_assert(W == 64 || H == 64, "sizes without a 64-point side use enc_tx_kernel");
  static constexpr int CW = W < 32 ? W : 32, CH = H < 32 ? H : 32;
  static constexpr int CODED = CW * CH;
  static constexpr int FR = CODED / W;  // dense forward rows inside the coded area
  static constexpr int THREADS = 128;
  // 64-wide blocks: the inverse row pass must have exactly one item per thread (load, barrier, store): THREADS / CH blocks.
  // 16 / 32-wide ones (no second layout: coded row r = dense row r, in place): one column per thread
  static constexpr bool TWO_LAYOUTS = W == 64;
  // GROUPS such sets of blocks per CTA: a slab pass is 9 000 - 13 600 instructions of straight-line code and ncu's top
  // stall was no_instruction (3.3 warps per issue at 64x64): with twice the blocks every fetched instruction does twice
  // the work, and the forward row pass (FR rows per block) has an item for every thread
  // (measured: 64x64 82 -> 74 us; the other four sizes lose 1 - 27 % with two groups -- shared memory then allows three
  // CTAs where 64x16 had four, and their transforms are shorter -- and keep one)
  static constexpr int GROUPS = (W == 64 && H == 64) ? ENC64_GROUPS : 1;
  static constexpr int BPG = THREADS / (TWO_LAYOUTS ? CH : W);  // blocks per group
  static constexpr int BPC = GROUPS * BPG;
#ifndef ENC64_MINB
#define ENC64_MINB 3
#endif
  static constexpr int MINB = H == 16 ? 4 : ENC64_MINB;  // CTAs per SM asked of ptxas (168 / 128 registers)
  static constexpr int STRIDE = W + 1;
  static constexpr int ROWS = FR > CH ? FR : CH;
  static constexpr int BLK = ROWS * STRIDE;
  static constexpr int LPB = 32, NCH = CODED / 32;
  static constexpr int LOGW = W == 16 ? 4 : W == 32 ? 5 : 6;
  static constexpr size_t SMEM = (size_t)BPC * BLK * 4;  // dynamic: up to 66.5 KB, three CTAs per SM
  static_assert(3 * SMEM <= 220 * 1024, "three CTAs per SM");
};

template <int W, int H, int CT, int RT>
__global__ void __launch_bounds__(Enc64Cfg<W, H>::THREADS, Enc64Cfg<W, H>::MINB) enc_tx64_kernel(const __grid_constant__ EncArgs a) {
  using C = Enc64Cfg<W, H>;
  constexpr int LW = C::LOGW, LH = H == 16 ? 4 : H == 32 ? 5 : 6;
  constexpr bool RECT = LW - LH == 1 || LH - LW == 1;  // get_rect_tx_log_ratio, inverse.rs:2070
  extern __shared__ __align__(16) int s_c[];  // [BPC][BLK]
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  const int nslabs = (int)((a.n + C::BPC - 1) / C::BPC);
  const int lts = (int)a.q.qc.log_tx_scale;
  for (int slab = blockIdx.x; slab < nslabs; slab += gridDim.x) {
    const long long blk0 = (long long)slab * C::BPC;
    const int nb = (int)min((long long)C::BPC, a.n - blk0);

    // ---- 1: diff + forward columns (forward.rs:1826-1869); rows >= FR of the result are never read ----
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
      int *dst = s_c + b * C::BLK + (a.row_flip ? (W - 1 - c) : c);
      if (a.rs1) {
#pragma unroll
        for (int r = 0; r < C::FR; r++) dst[r * C::STRIDE] = (int)((unsigned)o[r] + (unsigned)a.rnd1) >> a.rs1;
      } else {
#pragma unroll
        for (int r = 0; r < C::FR; r++) dst[r * C::STRIDE] = o[r];
      }
    }
    __syncthreads();

    // ---- 2: forward rows (forward.rs:1872-1898), in place, the FR dense rows of the coded area ------
#pragma unroll 1
    for (int it = tid; it < nb * C::FR; it += C::THREADS) {
      int *row = s_c + (it / C::FR) * C::BLK + (it % C::FR) * C::STRIDE;
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
    // scan positions index the first CODED words of the dense buffer: word p sits at p + (p >> log2 W) here
    for (int b = wrp; b < nb; b += C::THREADS / 32) {
      unsigned long long dist = quant::quantize_lanes<C::LPB, C::NCH, C::LOGW>(s_c + b * C::BLK, lane, 0, a.q, a.tx_dist != nullptr);
      if (a.tx_dist) {
#pragma unroll
        for (int m = 16; m; m >>= 1) dist += __shfl_xor_sync(0xffffffffu, dist, m);
        if (lane == 0) a.tx_dist[blk0 + b] = (dist + (1ull << (a.dist_bits - 1))) >> a.dist_bits;
      }
    }
    __syncthreads();

    // ---- 4a: quantized coefficients out (CODED per block), coalesced -------------------------------
    {
      int32_t *out = a.qcoeffs + blk0 * C::CODED;
      for (int i = tid; i < nb * C::CODED; i += C::THREADS) {
        const int b = i / C::CODED, p = i % C::CODED;
        out[i] = s_c[b * C::BLK + (p >> C::LOGW) * C::STRIDE + (p & (W - 1))];
      }
    }
    if (!a.reconstruct) {
      __syncthreads();  // s_c is rewritten by the next slab's column pass
      continue;
    }
    if (!C::TWO_LAYOUTS) __syncthreads();  // a row is read above by any thread and rewritten below by its own

    // ---- 4b: dequantize (quantize.rs:318-334) + inverse rows (inverse.rs:2073-2092) -----------------
    // coded row r = words [r * CW, (r + 1) * CW) of the buffer; its W outputs go to row r of the intermediate, on top
    // of other rows' inputs: every thread loads its row, then a barrier, then the stores
    auto inv_row = [&](int it) {
      const bool live = it < nb * C::CH;
      const int b = it / C::CH, r = it % C::CH;
      int v[W], o[W], l1 = 0;
      if (live) {
        const int *blk = s_c + b * C::BLK;
        const unsigned off = (1u << lts) - 1u;
#pragma unroll
        for (int c = 0; c < C::CW; c++) {
          const int p = r * C::CW + c;
          const int qv = blk[(p >> C::LOGW) * C::STRIDE + (p & (W - 1))];
          int t = (int)((unsigned)qv * ((c == 0 && r == 0) ? a.q.qc.dc_quant : a.q.qc.ac_quant) + ((unsigned)(qv >> 31) & off)) >> lts;
          if (RECT) t = (int)((unsigned)t * 2896u + 2048u) >> 12;  // round_shift(raw * INV_SQRT2, SQRT2_BITS)
          v[c] = t;
          l1 = (int)min((unsigned)l1 + (unsigned)abs(t), 0x40000000u);
        }
#pragma unroll
        for (int c = C::CW; c < W; c++) v[c] = 0;
      }
      if (C::TWO_LAYOUTS) __syncthreads();  // (every thread, once per group) also orders 4a's reads before the stores below
      if (live) {
        if (__all_sync(__activemask(), l1 + itx::NOCLAMP_MARGIN <= a.rg_row.hi)) {
          itx::inv1d<RT, W, false>(v, o, a.rg_row);
        } else {
#pragma unroll
          for (int c = 0; c < C::CW; c++) v[c] = itx::cl(v[c], a.rg_row);
          itx::inv1d<RT, W, true>(v, o, a.rg_row);
        }
        int *row = s_c + b * C::BLK + r * C::STRIDE;
        if (a.row_flip) {
#pragma unroll
          for (int c = 0; c < W; c++) row[W - 1 - c] = itx::cl((o[c] + a.irnd) >> a.ish, a.rg_col);
        } else {
#pragma unroll
          for (int c = 0; c < W; c++) row[c] = itx::cl((o[c] + a.irnd) >> a.ish, a.rg_col);
        }
      }
    };
    if constexpr (C::TWO_LAYOUTS) {
      // one group of whole blocks at a time (the hazard is inside a block): load, barrier, store
#pragma unroll 1
      for (int g = 0; g < C::GROUPS; g++) inv_row(tid + g * C::THREADS);
    } else {
#pragma unroll 1
      for (int it = tid; it < nb * C::CH; it += C::THREADS) inv_row(it);
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
        v[r] = r < C::CH ? col[r * C::STRIDE] : 0;
        if (r < C::CH) l1 += abs(v[r]);  // |v| < 2^17 after the row pass' clamp
      }
      if (__all_sync(__activemask(), l1 + itx::NOCLAMP_MARGIN <= a.rg_col.hi)) itx::inv1d<CT, H, false>(v, o, a.rg_col);
      else itx::inv1d<CT, H, true>(v, o, a.rg_col);
      constexpr int PXB = H < 32 ? H : 32;  // destination pixels in batches of independent loads
      if (a.px_bytes == 1) {
        uint8_t *p = (uint8_t *)a.rec.p0 + rec_off;
#pragma unroll
        for (int r0 = 0; r0 < H; r0 += PXB) {
          int px[PXB];
#pragma unroll
          for (int r = 0; r < PXB; r++) px[r] = p[(long long)(r0 + r) * a.rec.stride];
#pragma unroll
          for (int r = 0; r < PXB; r++) p[(long long)(r0 + r) * a.rec.stride] = (uint8_t)min(max(px[r] + (((a.col_flip ? o[H - 1 - (r0 + r)] : o[r0 + r]) + 8) >> 4), 0), a.maxv);
        }
      } else {
        uint16_t *p = (uint16_t *)a.rec.p0 + rec_off;
#pragma unroll
        for (int r0 = 0; r0 < H; r0 += PXB) {
          int px[PXB];
#pragma unroll
          for (int r = 0; r < PXB; r++) px[r] = p[(long long)(r0 + r) * a.rec.stride];
#pragma unroll
          for (int r = 0; r < PXB; r++) p[(long long)(r0 + r) * a.rec.stride] = (uint16_t)min(max(px[r] + (((a.col_flip ? o[H - 1 - (r0 + r)] : o[r0 + r]) + 8) >> 4), 0), a.maxv);
        }
      }
    }
    __syncthreads();  // s_c is rewritten by the next slab's column pass
  }
}

template <int W, int H, int CT, int RT>
int enc64_launch(r1c_ctx *ctx, const EncArgs &a) {
  using C = Enc64Cfg<W, H>;
  if (a.n == 0) return R1C_OK;
  // the attribute belongs to the (kernel, device) pair and the process may hold contexts on several devices: set it on
  // every launch, like me_sb2.cu does (the caller has made ctx->device current)
  cudaError_t e = cudaFuncSetAttribute(enc_tx64_kernel<W, H, CT, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
  if (e != cudaSuccess) return r1c_set_err(ctx, R1C_ERR_CUDA, "enc_tx64 kernel: %s", cudaGetErrorString(e));
  static int per_sm = 0;  // resident CTAs per SM of this instantiation
  if (!per_sm) {
    int nb = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, enc_tx64_kernel<W, H, CT, RT>, C::THREADS, C::SMEM);
    if (e != cudaSuccess || nb < 1) return r1c_set_err(ctx, R1C_ERR_CUDA, "enc_tx64 kernel does not fit an SM: %s", cudaGetErrorString(e));
    per_sm = nb;
  }
  const long long nslabs = (a.n + C::BPC - 1) / C::BPC;
  const long long full = (long long)ctx->sm_count * per_sm;
  const unsigned grid = (unsigned)(nslabs < full ? nslabs : full);
  enc_tx64_kernel<W, H, CT, RT><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

// per-size table: [ct*3 + rt] -> launcher, or nullptr: a 64-point side is a DCT, ADST goes up to 16 points (forward.rs:1762-1768,
// inverse.rs:1610-1634)
#define ENC64_BAD(T, N) (((T) == 1 && (N) > 16) || ((T) != 0 && (N) == 64))
#define ENC64_ENTRY(W, H, CT, RT)                                    \
  ((ENC64_BAD(CT, H) || ENC64_BAD(RT, W)) ? (enc_launch_fn) nullptr \
                                          : (enc_launch_fn)enc64_launch<W, H, ENC64_BAD(CT, H) ? 0 : (CT), ENC64_BAD(RT, W) ? 0 : (RT)>)

#define ENC64_SIZE_TABLE(NAME, W, H)                                                                          \
  extern const enc_launch_fn NAME[9];                                                                         \
  const enc_launch_fn NAME[9] = {ENC64_ENTRY(W, H, 0, 0), ENC64_ENTRY(W, H, 0, 1), ENC64_ENTRY(W, H, 0, 2),   \
                                 ENC64_ENTRY(W, H, 1, 0), ENC64_ENTRY(W, H, 1, 1), ENC64_ENTRY(W, H, 1, 2),   \
                                 ENC64_ENTRY(W, H, 2, 0), ENC64_ENTRY(W, H, 2, 1), ENC64_ENTRY(W, H, 2, 2)};

}  // namespace enc
