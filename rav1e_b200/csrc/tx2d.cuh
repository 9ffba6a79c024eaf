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
  // the same three entries as branch-free shift amounts (filled by the launcher): stage 0 is always a
  // left shift, stage 1 always a rounding right shift, stage 2 either
  int ls0, rs1, rnd1, ls2, rs2, rnd2;
  int col_flip, row_flip;
  // fused diff: residual = input - pred
  PlaneView input, pred;
  const r1c_tx_job *jobs;
  int bulk;                 // dense residual input is 16-byte aligned: bulk copies / 16-byte loads (set by the launcher)
  int out16;                // coefficient output is 16-byte aligned: 16-byte stores
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
  static constexpr int MAXD = W > H ? W : H, MIND = W < H ? W : H;
  // 256 threads for the small transforms, 128 for 32/64-point ones (their register footprint is large).
  // Blocks per CTA: THREADS / MAXD -- one column and one row per thread at most -- except for the 4:1 shapes
  // (4x16, 16x4, 8x32, 32x8, 16x64, 64x16), where the pass with the long transforms would run on a quarter of the
  // threads (barrier was the top stall, profiles/ncu_fht_r02.md): those take THREADS / MIND blocks, the long pass has one
  // item per thread and the short one loops
  static constexpr int THREADS = MAXD >= 32 ? 128 : 256;
  static constexpr int BPC = THREADS / (MAXD >= 4 * MIND ? MIND : MAXD);
  static constexpr int MID_STRIDE = W + 1;
  // residual blocks of narrow transforms sit 16 / 32 bytes apart in shared memory so that the lanes of a
  // warp that read the same (row, column) of different blocks hit different banks
  static constexpr int IN_PAD = W * H < 64 ? 0 : (W <= 8 ? 16 : (W == 16 ? 32 : 0));  // per-block copies of >= 128 B only
  static constexpr int IN_BLK = W * H * 2 + IN_PAD;  // bytes per staged block (multiple of 16)
  static constexpr int IN_BYTES = BPC * IN_BLK;     // one residual slab
  // two residual buffers (the next slab streams in while this one is transformed),
  // the padded i32 work array, two mbarriers
  static constexpr size_t SMEM = 2 * (size_t)IN_BYTES + (size_t)BPC * H * MID_STRIDE * 4 + 16;
};

// ---- async bulk copy (TMA 1-D) + mbarrier, PTX ISA 8.x ---------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(b))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "R1C_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra R1C_DONE;\n"
      "bra R1C_WAIT;\n"
      "R1C_DONE:\n"
      "}" ::"r"(smem_u32(b)),
      "r"(parity)
      : "memory");
}

// Persistent kernel: CTA c transforms slabs c, c + grid, c + 2 grid, ...
//
// The passes are issue-bound (ncu, 8x8 DCT: ~21 integer instructions per coefficient in
// the two 1-D transforms alone, 72-80 % of the issue slots busy), so everything around the
// butterflies is kept small:
//  * staging of dense residual input (FUSED_PB == 0), always one slab ahead of the math:
//      - W >= 32 and the smallest sizes: ONE cp.async.bulk per slab into the other half of a
//        double buffer, completion on an mbarrier that a single thread polls;
//      - 8- and 16-wide blocks: 16-byte loads into registers one slab ahead, stored to
//        shared memory with a 16 / 32-byte gap after every block, which removes the 4-way
//        bank conflicts of the column pass (blocks are 128 B apart otherwise);
//  * shifts with uniform amounts, flips hoisted into uniform branches;
//  * W <= 8: the row pass writes its W coefficients straight to global memory as 16-byte
//    stores (a thread owns W*4 contiguous bytes); wider rows go back through shared memory
//    for fully coalesced 4-byte stores.
template <int W, int H, int CT, int RT, int FUSED_PB>
__global__ void __launch_bounds__(FhtCfg<W, H>::THREADS) fht_kernel(FhtArgs a) {
  using C = FhtCfg<W, H>;
  constexpr int AREA = W * H;
  constexpr int BLK16 = C::IN_BLK / 2;          // staged block pitch in i16
  constexpr bool PREFETCH_REGS = C::IN_PAD > 0;  // else bulk copy (dense blocks)
  constexpr bool DIRECT_OUT = W <= 8;  // 16-wide rows: half-filled sectors per store instruction cost more than the detour
  constexpr int VPB = AREA * 2 / 16;                                   // 16-byte vectors per block
  constexpr int NV = (C::BPC * VPB + C::THREADS - 1) / C::THREADS;     // vectors per thread and slab
  extern __shared__ __align__(128) unsigned char smem_raw[];
  int16_t *s_in0 = (int16_t *)smem_raw;
  int *s_mid = (int *)(smem_raw + 2 * (size_t)C::IN_BYTES);
  uint64_t *mbar = (uint64_t *)(smem_raw + 2 * (size_t)C::IN_BYTES + (size_t)C::BPC * H * C::MID_STRIDE * 4);

  const int tid = threadIdx.x;
  const int nslabs = (int)((a.n + C::BPC - 1) / C::BPC);
  const bool fast = FUSED_PB == 0 && a.bulk;  // 16-byte aligned dense input
  const bool bulk = fast && !PREFETCH_REGS;
  const bool pref = fast && PREFETCH_REGS;
  int4 pre[NV];

  auto blocks_of = [&](int slab) { return (int)min((long long)C::BPC, a.n - (long long)slab * C::BPC); };
  auto issue_bulk = [&](int slab, int buf) {  // thread 0 only
    const uint32_t bytes = (uint32_t)blocks_of(slab) * AREA * 2;
    mbar_expect_tx(&mbar[buf], bytes);
    bulk_g2s(s_in0 + (size_t)buf * (C::IN_BYTES / 2), a.residual + (long long)slab * C::BPC * AREA, bytes, &mbar[buf]);
  };
  auto load_regs = [&](int slab) {
    const int4 *src = (const int4 *)(a.residual + (long long)slab * C::BPC * AREA);
    const int nv = blocks_of(slab) * VPB;
#pragma unroll
    for (int i = 0; i < NV; i++) {
      const int v = tid + i * C::THREADS;
      if (v < nv) pre[i] = __ldg(src + v);
    }
  };

  if (bulk) {
    if (tid == 0) {
      mbar_init(&mbar[0], 1);
      mbar_init(&mbar[1], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && (int)blockIdx.x < nslabs) issue_bulk(blockIdx.x, 0);
  } else if (pref) {
    if ((int)blockIdx.x < nslabs) load_regs(blockIdx.x);
  }

  int k = 0;
  for (int slab = blockIdx.x; slab < nslabs; slab += gridDim.x, k++) {
    const long long blk0 = (long long)slab * C::BPC;
    const int nb = blocks_of(slab);
    const int nxt = slab + gridDim.x;
    const int16_t *s_in = bulk ? s_in0 + (size_t)(k & 1) * (C::IN_BYTES / 2) : s_in0;

    // ---- stage the residual slab -------------------------------------------------
    if (bulk) {
      if (tid == 0) {
        // the other buffer was last read in the column pass of iteration k-1 (barriers ago)
        if (nxt < nslabs) issue_bulk(nxt, (k + 1) & 1);
        mbar_wait(&mbar[k & 1], (uint32_t)((k >> 1) & 1));
      }
    } else if (pref) {
      const int nv = nb * VPB;
#pragma unroll
      for (int i = 0; i < NV; i++) {
        const int v = tid + i * C::THREADS;
        if (v < nv) *(int4 *)((unsigned char *)s_in0 + (v / VPB) * C::IN_BLK + (v % VPB) * 16) = pre[i];
      }
      if (nxt < nslabs) load_regs(nxt);
    } else if (FUSED_PB == 0) {
      const int16_t *src = a.residual + blk0 * AREA;
      for (int i = tid; i < nb * AREA; i += C::THREADS) s_in0[(i / AREA) * BLK16 + (i % AREA)] = src[i];
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
        s_in0[b * BLK16 + rc] = (int16_t)(x - y);
      }
    }
    __syncthreads();

    // ---- columns (forward.rs:1826-1869) ----------------------------------------
    for (int itc = tid; itc < nb * W; itc += C::THREADS) {
      const int b = itc / W, c = itc % W;
      const int16_t *col = s_in + b * BLK16 + c;
      int v[H], o[H];
      if (a.col_flip) {  // :1829-1834 flip upside down
#pragma unroll
        for (int r = 0; r < H; r++) v[r] = (int)((unsigned)(int)col[(H - 1 - r) * W] << a.ls0);
      } else {
#pragma unroll
        for (int r = 0; r < H; r++) v[r] = (int)((unsigned)(int)col[r * W] << a.ls0);
      }
      fwd1d<CT, H>(v, o);
      const int cc = a.row_flip ? (W - 1 - c) : c;  // intended semantic of :1858-1862
      int *dst = s_mid + b * H * C::MID_STRIDE + cc;
      if (a.rs1) {
#pragma unroll
        for (int r = 0; r < H; r++) dst[r * C::MID_STRIDE] = (int)((unsigned)o[r] + (unsigned)a.rnd1) >> a.rs1;
      } else {
#pragma unroll
        for (int r = 0; r < H; r++) dst[r * C::MID_STRIDE] = o[r];
      }
    }
    __syncthreads();

    // ---- rows (forward.rs:1872-1898) ---------------------------------------------
    int32_t *out = a.coeffs + blk0 * AREA;
    const bool direct = DIRECT_OUT && a.out16;
    for (int itr = tid; itr < nb * H; itr += C::THREADS) {
      int *row = s_mid + itr * C::MID_STRIDE;  // itr = b*H + r
      int v[W], o[W];
#pragma unroll
      for (int c = 0; c < W; c++) v[c] = row[c];
      fwd1d<RT, W>(v, o);
      if (a.rs2) {
#pragma unroll
        for (int c = 0; c < W; c++) o[c] = (int)((unsigned)o[c] + (unsigned)a.rnd2) >> a.rs2;
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) o[c] = (int)((unsigned)o[c] << a.ls2);
      }
      if (direct) {
        int4 *d4 = (int4 *)(out + (long long)itr * W);
#pragma unroll
        for (int c = 0; c < W; c += 4) d4[c / 4] = make_int4(o[c], o[c + 1], o[c + 2], o[c + 3]);
      } else {
#pragma unroll
        for (int c = 0; c < W; c++) row[c] = o[c];
      }
    }
    __syncthreads();  // s_mid is rewritten by the next slab's column pass

    // ---- coalesced store of the coefficient slab (wide rows) --------------------
    if (!direct) {
      if (nb == C::BPC) {
        constexpr int ITER = C::BPC * AREA / C::THREADS;  // exact: THREADS divides the slab
        constexpr int ROWS_PER_ITER = C::THREADS / W;     // W divides THREADS
        const int *src = s_mid + (tid / W) * C::MID_STRIDE + (tid % W);
#pragma unroll
        for (int it = 0; it < ITER; it++) out[it * C::THREADS + tid] = src[it * ROWS_PER_ITER * C::MID_STRIDE];
      } else {
        for (int i = tid; i < nb * AREA; i += C::THREADS) out[i] = s_mid[(i / W) * C::MID_STRIDE + (i % W)];
      }
      __syncthreads();
    }
  }
}

typedef int (*fht_launch_fn)(r1c_ctx *, const FhtArgs &, int fused_pb);

template <int W, int H, int CT, int RT, int PBK>
int fht_launch_one(r1c_ctx *ctx, FhtArgs a) {
  using C = FhtCfg<W, H>;
  static int per_sm = 0;  // resident CTAs per SM of this instantiation
  if (!per_sm) {
    cudaError_t e = cudaFuncSetAttribute(fht_kernel<W, H, CT, RT, PBK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM);
    if (e != cudaSuccess) return r1c_set_err(ctx, R1C_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int nb = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fht_kernel<W, H, CT, RT, PBK>, C::THREADS, C::SMEM);
    if (e != cudaSuccess || nb < 1) return r1c_set_err(ctx, R1C_ERR_CUDA, "fht kernel does not fit an SM: %s", cudaGetErrorString(e));
    per_sm = nb;
  }
  const long long nslabs = (a.n + C::BPC - 1) / C::BPC;
  const long long full = (long long)ctx->sm_count * per_sm;  // persistent: exactly the resident CTAs
  const unsigned grid = (unsigned)(nslabs < full ? nslabs : full);
  // bulk copies need 16-byte aligned global addresses (slab sizes are multiples of 32 bytes)
  a.bulk = PBK == 0 && (((uintptr_t)a.residual) & 15) == 0;
  a.out16 = (((uintptr_t)a.coeffs) & 15) == 0;
  fht_kernel<W, H, CT, RT, PBK><<<grid, C::THREADS, C::SMEM, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

template <int W, int H, int CT, int RT>
int fht_launch(r1c_ctx *ctx, const FhtArgs &a, int fused_pb) {
  if (a.n == 0) return R1C_OK;
  if (fused_pb == 0) return fht_launch_one<W, H, CT, RT, 0>(ctx, a);
  if (fused_pb == 1) return fht_launch_one<W, H, CT, RT, 1>(ctx, a);
  return fht_launch_one<W, H, CT, RT, 2>(ctx, a);
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
