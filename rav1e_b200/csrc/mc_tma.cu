// mc_tma.cu -- put_8tap / prep_8tap (reference src/mc.rs:213-387) around source windows staged by 2-D TMA.
//
// The first-generation kernel (mc.cu) fetched every thread's source rows itself (six 4-byte cp.async + six LDS.32 per
// row for four output pixels, every lane in another 32-byte sector): ncu showed it bound by the load / store unit
// (mio_throttle 8 cycles per issue, 19 sectors per request), at 0.13 of the HBM roofline.  Here the LSU never touches
// global source pixels:
//
//  * unit = one job, or one 32x32 tile of a larger job; units are binned by (width class, height class) into 16
//    classes by two small kernels (count, scatter) -- no prefix sum over the job list;
//  * per class ONE tensor map of the source plane with box = (tile_w + 7 + alignment slack, in 16-byte steps) x (tile_h + 7);
//    a producer warp issues one cp.async.bulk.tensor.2d per unit (box origin = src - 3 floored to 16 bytes, out-of-plane
//    parts arrive as zeros and only ever meet zero taps) into a two-stage shared-memory ring, completion on an
//    mbarrier per stage; it also drops the unit's job record next to the pixels, so the compute warps read nothing
//    from global memory;
//  * compute: a thread owns a strip of 4 output columns and walks down the unit's rows.  The box starts at src - 3
//    rounded down to 16 bytes (a hardware rule of tensor loads), so pixel (x0 - 3) of a strip sits at a known phase
//    inside a word of the staged row: one PRMT (u16) / funnel shift (u8) per word realigns it.  Odd columns
//    use byte-shifted copies of the packed s8 taps instead of re-packed pixels (u8: 2 / 3 / 3 / 3 DP4A for the four
//    columns, u16: 4 / 5 / 4 / 5 DP2A), the horizontally filtered rows are kept as (even row, odd row) i16 pairs in
//    a five-deep register ring and every pair of output rows costs 4 + 5 DP2A per column.
//  * ONE formula for the four (col_frac, row_frac) cases, taps halved, exactly as in mc.cu (see its comment on
//    walk_strip_scalar): the result is bit-exact with put_8tap_ref / prep_8tap_ref.
// Bound: HBM by bytes (2 * px_bytes per output pixel), INT issue in practice (~13.5 thread instructions per output
// pixel for 32-row strips, mostly half-rate DP2A / DP4A); small blocks are bound by L2 -> SM sectors of their halo.
#include <cuda.h>

#include <mutex>

#include "common.cuh"
#include "mc_filters.cuh"

namespace {

constexpr int NCLS = 12;               // width class (w <= 4, 8, 16, 32) x height class (h <= 4, 8, 16)
constexpr int NHC = 3;
constexpr int CT = 128;                // compute threads per CTA
constexpr int THREADS = CT + 32;       // + the producer warp
constexpr int STAGE_BYTES = 36864;     // pixels of one stage (16 units of the largest box: 48 px x 24 rows of u16)
constexpr int MAX_UPB = 128;           // units per batch, at most
constexpr int REC_WORDS = 8;           // job record per unit: the 7 words of r1c_mc_job + (tile_x | tile_y << 8)
#ifndef MC_TMA_STAGES
#define MC_TMA_STAGES 1
#endif
constexpr int STAGES = MC_TMA_STAGES;  // 1: the CTAs of an SM (4 resident) overlap each other's loads; 2: double buffer inside the CTA

__host__ __device__ inline int wclass_of(int w) { return w <= 4 ? 0 : w <= 8 ? 1 : w <= 16 ? 2 : 3; }
__host__ __device__ inline int hclass_of(int h) { return h <= 4 ? 0 : h <= 8 ? 1 : 2; }
__host__ __device__ inline int strips_of(int wc) { return 1 << wc; }          // 4-column strips of a unit
__host__ __device__ inline int tile_h_of(int hc) { return 4 << hc; }
// staged row of a unit, in pixels.  The hardware wants the global address of a box row on a 16-byte boundary (a
// tensor load whose inner coordinate * element size is not a multiple of 16 faults with "illegal instruction";
// tools/tma_probe2.cu), so the box starts at (src_x - 3) rounded DOWN to 16 bytes and is up to 16 / pb - 1 pixels wider
// than tile width + 7; rounded up to 16 bytes.
#ifndef MC_TMA_W2PAD
#define MC_TMA_W2PAD 0  // 1: the 16-wide u16 box is 16 bytes wider (row pitch 20 words: four bank positions instead of two)
#endif
__host__ __device__ inline int box_w_of(int wc, int pb) {
  const int tw = wc == 0 ? 8 : (4 << wc);  // w <= 4 shares the 8-wide box
  const int px16 = 16 / pb;
  return (tw + 7 + (px16 - 1) + px16 - 1) / px16 * px16 + ((MC_TMA_W2PAD && wc == 2 && pb == 2) ? px16 : 0);
}
// Units of the two narrow width classes are owned by one or two threads, so the lanes of a warp read the same (row, word)
// of different unit slots, and slots are 128-byte aligned (what a tensor load demands: tools/tma_probe3.cu): every lane
// on one bank.  Their boxes start (unit & STAG) rows higher, which moves a lane's rows by that many row pitches
// (12 words for the 8-wide u16 box: 4 distinct bank groups).  The 16-wide class (4 strips per unit, row pitch 16 words)
// has two distinct positions: one extra row.  Measured: 8 phases lose (more bytes, fewer units per batch), and the 32-wide
// class (8 strips per unit, bound by the XU pipe) does not gain.
#ifndef MC_TMA_STAG
#define MC_TMA_STAG 3
#endif
__host__ __device__ inline int stag_of(int wc, int pb) {
  return wc <= 1 ? MC_TMA_STAG : (wc == 2 ? ((pb == 1 || MC_TMA_W2PAD) ? MC_TMA_STAG : 1) : 0);
}
__host__ __device__ inline int box_h_of(int wc, int hc, int pb) { return tile_h_of(hc) + 7 + stag_of(wc, pb); }
// + one row: the last (even, odd) pair of a strip reads one row past the box (it meets a zero tap)
__host__ __device__ inline int slot_bytes_of(int wc, int hc, int pb) {
  return (box_w_of(wc, pb) * pb * (box_h_of(wc, hc, pb) + 1) + 127) & ~127;
}
__host__ __device__ inline int upb_of(int wc, int hc, int pb) {
  const int a = CT / strips_of(wc), b = STAGE_BYTES / slot_bytes_of(wc, hc, pb);
  return a < b ? a : b;
}

struct TmaArgs {
  alignas(64) CUtensorMap map[9];   // [box width class 0..2][height class 0..2]
  PlaneView src;
  uint8_t *dst0;
  int dst_stride;
  int16_t *tmp;
  const r1c_mc_job *jobs;
  const uint32_t *units;       // unit list, class-major: job index | tile << 26
  const uint32_t *cls_count;   // [12] units per class (written by the count kernel)
  long long n;
  int bit_depth;
};

// ---- binning ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void unit_shape(const r1c_mc_job &j, int &cls, int &tiles_x, int &nunits) {
  tiles_x = (j.w + 31) >> 5;
  const int tiles_y = (j.h + 15) >> 4;
  nunits = tiles_x * tiles_y;  // <= 4 x 8
  const int w = j.w > 32 ? 32 : j.w, h = j.h > 16 ? 16 : j.h;  // edge tiles of large jobs may be smaller: same class, fewer live strips / rows
  cls = wclass_of(w) * NHC + hclass_of(h);
}

// pass 1: units per class.  cursor[0..15] = counts, then (pass 2) running offsets; counts[2 * NCLS] = number of jobs
// whose size differs from job 0's (or is empty): 0 = a list of ONE block size, the usual batch, for which unit u is tile
// u % units_per_job of job u / units_per_job -- the scatter pass returns at once and the main kernel needs no unit list
// (the three passes over the 44-byte job records were 1/6 of a call: a 4x4 job moves 64 bytes of pixels)
__global__ void mc_count_kernel(const r1c_mc_job *jobs, long long n, uint32_t *counts) {
  __shared__ uint32_t s_cnt[NCLS];
  __shared__ uint32_t s_odd;
  if (threadIdx.x < NCLS) s_cnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_odd = 0;
  __syncthreads();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const r1c_mc_job j = jobs[i];
    const uint32_t wh0 = __ldg((const uint32_t *)jobs + 4);  // w | h << 16 of job 0
    bool odd = true;
    if (j.w > 0 && j.h > 0) {
      int cls, tx, nu;
      unit_shape(j, cls, tx, nu);
      atomicAdd(&s_cnt[cls], (uint32_t)nu);
      odd = ((uint32_t)(uint16_t)j.w | ((uint32_t)(uint16_t)j.h << 16)) != wh0;
    }
    if (__any_sync(__activemask(), odd) && odd) s_odd = 1;
  }
  __syncthreads();
  if (threadIdx.x < NCLS && s_cnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], s_cnt[threadIdx.x]);
  if (threadIdx.x == 0 && s_odd) atomicAdd(&counts[2 * NCLS], 1u);
}

// pass 2: scatter (job, tile) into the class-major unit list.  counts[NCLS ..] = cursors (zeroed by the host memset)
__global__ void mc_scatter_kernel(const r1c_mc_job *jobs, long long n, uint32_t *counts, uint32_t *units) {
  if (counts[2 * NCLS] == 0) return;  // one block size: no unit list
  __shared__ uint32_t s_cnt[NCLS], s_base[NCLS];
  if (threadIdx.x < NCLS) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int cls = 0, tx = 1, nu = 0;
  uint32_t local = 0;
  if (i < n) {
    const r1c_mc_job j = jobs[i];
    if (j.w > 0 && j.h > 0) {
      unit_shape(j, cls, tx, nu);
      local = atomicAdd(&s_cnt[cls], (uint32_t)nu);
    }
  }
  __syncthreads();
  if (threadIdx.x < NCLS) {
    uint32_t start = 0;
    for (int c = 0; c < threadIdx.x; c++) start += counts[c];
    s_base[threadIdx.x] = start + (s_cnt[threadIdx.x] ? atomicAdd(&counts[NCLS + threadIdx.x], s_cnt[threadIdx.x]) : 0u);
  }
  __syncthreads();
  for (int u = 0; u < nu; u++) units[s_base[cls] + local + u] = (uint32_t)i | ((uint32_t)u << 26);
}

// ---- PTX helpers -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) {  // no arrival: the lane arrives after its stores
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MCT_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MCT_DONE;\n"
      "nanosleep.u32 40;\n"  // the spin loops of waiting warps cost issue slots of the working ones
      "bra MCT_WAIT;\n"
      "MCT_DONE:\n"
      "}" ::"r"(s32(b)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *b) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(dst)),
               "l"((unsigned long long)map), "r"(x), "r"(y), "r"(s32(b))
               : "memory");
}
__device__ __forceinline__ int dp4a_us(uint32_t px, uint32_t taps, int acc) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_lo_us(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.lo.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_hi_us(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.hi.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_lo_ss(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_hi_ss(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}

// get_filter, src/mc.rs:201-210
__device__ __forceinline__ int filter_bank(int mode, int length) {
  return (mode == R1C_BILINEAR || length > 4) ? mode : ((mode < 1 ? mode : 1) + 4);
}

// the 8 halved taps t0..t7 as s8, and their copies shifted by 1 / 2 / 3 bytes (zero-filled): column c of a strip
// whose pixels start c elements into the aligned words multiplies word k by shifted(c)[k]
struct Taps {
  uint32_t a0, a1;        // (t0 t1 t2 t3) (t4 t5 t6 t7)
  uint32_t b0, b1, b2;    // (0 t0 t1 t2) (t3 t4 t5 t6) (t7 0 0 0)
  __device__ __forceinline__ void set(uint32_t lo, uint32_t hi) {
    a0 = lo;
    a1 = hi;
    b0 = lo << 8;
    b1 = (lo >> 24) | (hi << 8);
    b2 = hi >> 24;
  }
};

// horizontal sums of the 4 columns of a strip from the staged row; `row` = the aligned word holding pixel x0 - 3,
// `phase` = its position inside that word (u8: shift in bits, u16: PRMT selector)
template <int PB>
__device__ __forceinline__ void hsum4(const uint32_t *row, uint32_t phase, const Taps &t, uint32_t c0, uint32_t c1, uint32_t c2,
                                      uint32_t d0, uint32_t d1, uint32_t d2, int rnd, int (&s)[4]) {
  if (PB == 1) {
    // bytes x0-3 .. x0+7 = words 0..2 after the funnel shift; column c needs bytes c .. c+7
    const uint32_t r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];
    const uint32_t w0 = __funnelshift_r(r0, r1, phase), w1 = __funnelshift_r(r1, r2, phase), w2 = __funnelshift_r(r2, r3, phase);
    s[0] = dp4a_us(w1, t.a1, dp4a_us(w0, t.a0, rnd));
    s[1] = dp4a_us(w2, t.b2, dp4a_us(w1, t.b1, dp4a_us(w0, t.b0, rnd)));
    s[2] = dp4a_us(w2, c2, dp4a_us(w1, c1, dp4a_us(w0, c0, rnd)));
    s[3] = dp4a_us(w2, d2, dp4a_us(w1, d1, dp4a_us(w0, d0, rnd)));
  } else {
    // pixels x0-3 .. x0+7 = halves of words 0..5 after the PRMT; column c needs pixels c .. c+7
    const uint32_t r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3], r4 = row[4], r5 = row[5], r6 = row[6];
    const uint32_t e0 = __byte_perm(r0, r1, phase), e1 = __byte_perm(r1, r2, phase), e2 = __byte_perm(r2, r3, phase),
                   e3 = __byte_perm(r3, r4, phase), e4 = __byte_perm(r4, r5, phase), e5 = __byte_perm(r5, r6, phase);
    s[0] = dp2a_hi_us(e3, t.a1, dp2a_lo_us(e2, t.a1, dp2a_hi_us(e1, t.a0, dp2a_lo_us(e0, t.a0, rnd))));
    s[1] = dp2a_lo_us(e4, t.b2, dp2a_hi_us(e3, t.b1, dp2a_lo_us(e2, t.b1, dp2a_hi_us(e1, t.b0, dp2a_lo_us(e0, t.b0, rnd)))));
    s[2] = dp2a_hi_us(e4, t.a1, dp2a_lo_us(e3, t.a1, dp2a_hi_us(e2, t.a0, dp2a_lo_us(e1, t.a0, rnd))));
    s[3] = dp2a_lo_us(e5, t.b2, dp2a_hi_us(e4, t.b1, dp2a_lo_us(e3, t.b1, dp2a_hi_us(e2, t.b0, dp2a_lo_us(e1, t.b0, rnd)))));
  }
}

template <int PB, int KIND>
__device__ __forceinline__ void store_row(uint8_t *dp, bool vec, int ncols, const int (&v)[4], uint32_t mx) {
  // every result fits i16 (a pixel plus filter overshoot; prep keeps the low 16 bits like `as i16`)
  uint32_t p01 = __byte_perm((uint32_t)v[0], (uint32_t)v[1], 0x5410), p23 = __byte_perm((uint32_t)v[2], (uint32_t)v[3], 0x5410);
  if (KIND != 1) {
    p01 = __vmins2(__vmaxs2(p01, 0u), mx);
    p23 = __vmins2(__vmaxs2(p23, 0u), mx);
  }
  if (vec) {
    if (KIND != 1 && PB == 1) *(uint32_t *)dp = __byte_perm(p01, p23, 0x6420);
    else *(uint2 *)dp = make_uint2(p01, p23);
  } else {
    const uint32_t q[2] = {p01, p23};
#pragma unroll
    for (int c = 0; c < 4; c++)
      if (c < ncols) {
        const uint32_t e = (q[c >> 1] >> (16 * (c & 1))) & 0xffffu;
        if (KIND != 1 && PB == 1) dp[c] = (uint8_t)e;
        else ((uint16_t *)dp)[c] = (uint16_t)e;
      }
  }
}

// KIND 0 = put (clamped pixels into the dst plane), 1 = prep (i16, dense w*h at tmp_off), 2 = put into a dense w*h
// pixel block at tmp_off (sub-pel search)
template <int PB, int KIND>
__global__ void __launch_bounds__(THREADS, STAGES == 1 ? 5 : 2) mc_tma_kernel(const __grid_constant__ TmaArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  // [stage pixels x 2][records x 2][mbarriers full x 2, empty x 2]
  unsigned char *s_px = smem;
  uint32_t *s_rec = (uint32_t *)(smem + STAGES * STAGE_BYTES);
  uint64_t *s_full = (uint64_t *)(s_rec + STAGES * MAX_UPB * REC_WORDS);
  uint64_t *s_empty = s_full + STAGES;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&s_full[s], 32);  // the 32 producer lanes + the bytes of their TMA loads
      mbar_init(&s_empty[s], CT / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;

  // the same batch sequence in the producer and the consumers: classes in order, batches b = blockIdx.x, + gridDim.x ...
  struct Cursor {
    int cls;
    uint32_t b, start, cnt, nbatch;
    int wc, hc, upb;
    __device__ __forceinline__ void next_class(const TmaArgs &a) {
      for (;;) {
        start += cnt;
        cls++;
        if (cls >= NCLS) {
          cnt = 0;
          return;
        }
        cnt = __ldg(&a.cls_count[cls]);
        if (cnt == 0) continue;
        wc = cls / NHC;
        hc = cls - wc * NHC;
        upb = upb_of(wc, hc, PB);
        nbatch = (cnt + upb - 1) / upb;
        if (blockIdx.x < nbatch) {
          b = blockIdx.x;
          return;
        }
      }
    }
    __device__ __forceinline__ void init(const TmaArgs &a) {
      cls = -1;
      start = cnt = 0;
      next_class(a);
    }
    __device__ __forceinline__ void advance(const TmaArgs &a) {
      b += gridDim.x;
      if (b >= nbatch) next_class(a);
    }
    __device__ __forceinline__ bool done() const { return cls >= NCLS; }
  };
  Cursor cur;
  cur.init(a);
  uint32_t it = 0;  // batches of this CTA so far: stage = it % STAGES, phase = (it / STAGES) & 1
  if (warp == CT / 32) {
    // ================= producer warp =================
    // The unit and job records of a batch are fetched one batch ahead (while the compute warps work on the previous
    // one), so that only the tensor loads themselves lie between "stage free" and "stage full".
    constexpr int UPL = MAX_UPB / 32;  // units per lane
    uint32_t pw[UPL][7];  // [0] = tile index, [1..6] = words 1..6 of r1c_mc_job
    uint32_t pdx[UPL];
    // a list of one block size (mc_count_kernel): unit U is tile U % npj of job U / npj, there is no unit list
    const bool uni = __ldg(&a.cls_count[2 * NCLS]) == 0;
    uint32_t npj = 1;
    if (uni) {
      int c0, t0, n0;
      unit_shape(a.jobs[0], c0, t0, n0);
      npj = (uint32_t)n0;
    }
    auto fetch = [&](const Cursor &c) {
      const uint32_t u0 = c.b * c.upb, nu = min((uint32_t)c.upb, c.cnt - u0);
#pragma unroll
      for (int k = 0; k < UPL; k++) {
        const uint32_t u = lane + 32 * k;
        if (u < nu) {
          uint32_t e;
          if (uni) {
            const uint32_t uu = c.start + u0 + u, jb = uu / npj;
            e = jb | ((uu - jb * npj) << 26);
          } else {
            e = __ldg(&a.units[c.start + u0 + u]);
          }
          const uint32_t *jw = (const uint32_t *)(a.jobs + (e & 0x3ffffffu));
          pdx[k] = __ldg(jw);
#pragma unroll
          for (int i = 1; i < 7; i++) pw[k][i] = __ldg(jw + i);
          pw[k][0] = e >> 26;
        }
      }
    };
    if (!cur.done()) fetch(cur);
    while (!cur.done()) {
      const int st = it % STAGES;
      const int wc = cur.wc, hc = cur.hc;
      const int slot = slot_bytes_of(wc, hc, PB);
      const CUtensorMap *map = &a.map[(wc == 0 ? 0 : wc - 1) * NHC + hc];
      if (it >= STAGES) mbar_wait(&s_empty[st], ((it / STAGES) - 1) & 1);
      const uint32_t u0 = cur.b * cur.upb, nu = min((uint32_t)cur.upb, cur.cnt - u0);
      if (lane == 0) mbar_expect_tx(&s_full[st], nu * (uint32_t)(box_w_of(wc, PB) * PB * box_h_of(wc, hc, PB)));
#pragma unroll
      for (int k = 0; k < UPL; k++) {
        const uint32_t u = lane + 32 * k;
        if (u < nu) {
          uint32_t *rec = s_rec + (st * MAX_UPB + u) * REC_WORDS;
          const uint32_t tile = pw[k][0];
          const int jwid = (int)(int16_t)(pw[k][4] & 0xffffu);
          const int tiles_x = (jwid + 31) >> 5;
          const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
          rec[0] = pdx[k];
#pragma unroll
          for (int i = 1; i < 7; i++) rec[i] = pw[k][i];
          const int x = a.src.xorigin + (int)pw[k][2] + 32 * tx - 3, y = a.src.yorigin + (int)pw[k][3] + 16 * ty - 3;
          const int x_al = x & ~(16 / PB - 1);  // floor to 16 bytes (two's complement: also for negative x)
          rec[7] = (uint32_t)tx | ((uint32_t)ty << 8) | ((uint32_t)(x - x_al) << 16);
          tma_load_2d(s_px + st * STAGE_BYTES + u * slot, map, x_al, y - (int)(u & (uint32_t)stag_of(wc, PB)), &s_full[st]);
        }
      }
      mbar_arrive(&s_full[st]);  // release: after this lane's record stores
      cur.advance(a);
      if (!cur.done()) fetch(cur);
      it++;
    }
  } else {
    while (!cur.done()) {
      const int wc = cur.wc, hc = cur.hc;
      const int upb = cur.upb;
      const int slot = slot_bytes_of(wc, hc, PB);
      const int pitch = box_w_of(wc, PB) * PB;  // bytes per staged row
      const uint32_t cnt = cur.cnt;
      {
      // ================= compute warps =================
      const int strips = strips_of(wc);
      const int un = tid >> wc, strip = tid & (strips - 1);
      {
        const int st = it % STAGES;
        const uint32_t u0 = cur.b * upb, nu = min((uint32_t)upb, cnt - u0);
        mbar_wait(&s_full[st], (it / STAGES) & 1);
        if ((uint32_t)un < nu) {
          const uint32_t *rec = s_rec + (st * MAX_UPB + un) * REC_WORDS;
          const int dst_x = (int)rec[0], dst_y = (int)rec[1];
          const int jw = (int)(int16_t)(rec[4] & 0xffffu), jh = (int)(int16_t)(rec[4] >> 16);
          const int col_frac = (int)(int8_t)(rec[5] & 0xff), row_frac = (int)(int8_t)((rec[5] >> 8) & 0xff);
          const int mode_x = (int)(int8_t)((rec[5] >> 16) & 0xff), mode_y = (int)(int8_t)(rec[5] >> 24);
          const long long tmp_off = (long long)(int)rec[6];
          const int tx = (int)(rec[7] & 0xff), ty = (int)((rec[7] >> 8) & 0xff), dx = (int)(rec[7] >> 16);
          const int x0 = 32 * tx + 4 * strip;          // first column of the strip inside the job
          const int ncols = min(4, jw - x0);
          const int r0 = 16 * ty, nout = min(jh - r0, tile_h_of(hc));
          if (ncols > 0 && nout > 0) {
            Taps hx, vy;
            {
              const int bx = filter_bank(mode_x, jw) * 16 + (col_frac & 15), by = filter_bank(mode_y, jh) * 16 + (row_frac & 15);
              hx.set(__ldg(&g_half_taps.lo[bx]), __ldg(&g_half_taps.hi[bx]));
              vy.set(__ldg(&g_half_taps.lo[by]), __ldg(&g_half_taps.hi[by]));
            }
            // u8 rows: the 2- and 3-byte shifted copies as well
            const uint32_t c0 = hx.a0 << 16, c1 = (hx.a0 >> 16) | (hx.a1 << 16), c2 = hx.a1 >> 16;
            const uint32_t d0 = hx.a0 << 24, d1 = (hx.a0 >> 8) | (hx.a1 << 24), d2 = hx.a1 >> 8;
            const int hsh = 6 - ib, hrnd = 1 << (5 - ib);
            const int vshift = (KIND == 1 ? 7 : 7 + ib) - 1, vrnd = 1 << (vshift - 1);
            const uint32_t mx = (uint32_t)maxv * 0x10001u;
            // pixel (x0 - 3) of the strip sits dx + 4 * strip pixels into the staged row: word-aligned base + a phase of
            // 0 / 1 pixel (u16) or 0..3 bytes (u8), undone per row by one PRMT / funnel shift per word
            const int boff = (dx + 4 * strip) * PB;
            const uint32_t phase = PB == 1 ? (uint32_t)(boff & 3) * 8u : ((boff & 2) ? 0x5432u : 0x3210u);
            const unsigned char *rowp = s_px + st * STAGE_BYTES + un * slot + (un & stag_of(wc, PB)) * pitch + (boff & ~3);
            uint8_t *dp;
            long long dstride;
            constexpr int OB = KIND == 1 ? 2 : PB;  // bytes per output element
            if (KIND == 0) {
              dp = a.dst0 + ((long long)(dst_y + r0) * a.dst_stride + dst_x + x0) * PB;
              dstride = (long long)a.dst_stride * PB;
            } else {
              dp = (uint8_t *)a.tmp + (tmp_off + (long long)r0 * jw + x0) * OB;
              dstride = (long long)jw * OB;
            }
            const bool vec = ncols == 4 && (((uintptr_t)dp | (uintptr_t)dstride) & (uintptr_t)(4 * OB - 1)) == 0;
            // Filtered rows in (even, odd) pairs: pair p = rows 2p, 2p + 1 of the box (row i = source row r0 - 3 + i).
            // Output rows 2q and 2q + 1 need filtered rows 2q .. 2q + 8 = pairs q .. q + 4: the even one with the taps as
            // they are on pairs q .. q + 3, the odd one with the taps shifted by one row on pairs q .. q + 4.
            uint32_t ring[4][5];
            auto hpair = [&](int slot_) {
              int se[4], so[4];
              hsum4<PB>((const uint32_t *)rowp, phase, hx, c0, c1, c2, d0, d1, d2, hrnd, se);
              hsum4<PB>((const uint32_t *)(rowp + pitch), phase, hx, c0, c1, c2, d0, d1, d2, hrnd, so);
              rowp += 2 * pitch;
#pragma unroll
              for (int c = 0; c < 4; c++) ring[c][slot_] = __byte_perm((uint32_t)(se[c] >> hsh), (uint32_t)(so[c] >> hsh), 0x5410);
            };
            hpair(0);
            hpair(1);
            hpair(2);
            hpair(3);
            auto vpair = [&](int u, bool both) {  // output rows 2q, 2q + 1 from ring slots u .. u + 4 (mod 5)
              int v[4];
#pragma unroll
              for (int c = 0; c < 4; c++) {
                int t = dp2a_lo_ss(ring[c][u], vy.a0, vrnd);
                t = dp2a_hi_ss(ring[c][(u + 1) % 5], vy.a0, t);
                t = dp2a_lo_ss(ring[c][(u + 2) % 5], vy.a1, t);
                t = dp2a_hi_ss(ring[c][(u + 3) % 5], vy.a1, t);
                v[c] = t >> vshift;
              }
              store_row<PB, KIND>(dp, vec, ncols, v, mx);
              if (both) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                  int t = dp2a_lo_ss(ring[c][u], vy.b0, vrnd);
                  t = dp2a_hi_ss(ring[c][(u + 1) % 5], vy.b0, t);
                  t = dp2a_lo_ss(ring[c][(u + 2) % 5], vy.b1, t);
                  t = dp2a_hi_ss(ring[c][(u + 3) % 5], vy.b1, t);
                  t = dp2a_lo_ss(ring[c][(u + 4) % 5], vy.b2, t);
                  v[c] = t >> vshift;
                }
                store_row<PB, KIND>(dp + dstride, vec, ncols, v, mx);
              }
              dp += 2 * dstride;
            };
            if (nout == 16) {  // the common unit: a full-height tile, no per-row predicates
#pragma unroll
              for (int q = 0; q < 8; q++) {
                hpair((q + 4) % 5);
                vpair(q % 5, true);
              }
            } else {
              const int nq = (nout + 1) >> 1;
              for (int q0 = 0; q0 < nq; q0 += 5) {
#pragma unroll
                for (int u = 0; u < 5; u++) {
                  const int q = q0 + u;
                  if (q < nq) {
                    hpair((u + 4) % 5);
                    vpair(u, 2 * q + 1 < nout);
                  }
                }
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[st]);
      }
      }
      cur.advance(a);
      it++;
    }
  }
}

// ---- host ------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
int g_encode_state = 0;  // 0 = not looked up, 1 = ok, -1 = unavailable
std::mutex g_mu;

struct MapCacheEntry {
  const void *data;
  int64_t stride, alloc_h;
  int pb, device;
  CUtensorMap map[9];
};
constexpr int MAP_CACHE = 16;
MapCacheEntry g_cache[MAP_CACHE];
int g_cache_n = 0, g_cache_next = 0;

bool get_maps(const r1c_plane *src, int device, CUtensorMap *out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_encode_state == 0) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess && fn &&
        qres == cudaDriverEntryPointSuccess) {
      g_encode = (EncodeTiledFn)fn;
      g_encode_state = 1;
    } else {
      g_encode_state = -1;
    }
  }
  if (g_encode_state < 0) return false;
  for (int i = 0; i < g_cache_n; i++) {
    const MapCacheEntry &e = g_cache[i];
    if (e.data == src->d_data && e.stride == src->cfg.stride && e.alloc_h == src->cfg.alloc_height && e.pb == src->px_bytes &&
        e.device == device) {
      memcpy(out, e.map, sizeof(e.map));
      return true;
    }
  }
  MapCacheEntry e;
  e.data = src->d_data;
  e.stride = src->cfg.stride;
  e.alloc_h = src->cfg.alloc_height;
  e.pb = src->px_bytes;
  e.device = device;
  const cuuint64_t dims[2] = {(cuuint64_t)e.stride, (cuuint64_t)e.alloc_h};
  const cuuint64_t strides[1] = {(cuuint64_t)e.stride * (cuuint64_t)e.pb};
  const cuuint32_t estr[2] = {1, 1};
  for (int wc = 1; wc < 4; wc++)
    for (int hc = 0; hc < NHC; hc++) {
      const cuuint32_t box[2] = {(cuuint32_t)box_w_of(wc, e.pb), (cuuint32_t)box_h_of(wc, hc, e.pb)};
      const CUresult r = g_encode(&e.map[(wc - 1) * NHC + hc], e.pb == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2,
                                  const_cast<void *>(e.data), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) return false;
    }
  const int slot = g_cache_n < MAP_CACHE ? g_cache_n++ : (g_cache_next++ % MAP_CACHE);
  g_cache[slot] = e;
  memcpy(out, e.map, sizeof(e.map));
  return true;
}

constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + (size_t)STAGES * MAX_UPB * REC_WORDS * 4 + 2 * STAGES * 8;

template <int PB, int KIND>
int launch_kernel(r1c_ctx *ctx, const TmaArgs &a) {
  auto kern = mc_tma_kernel<PB, KIND>;
  R1C_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, THREADS, SMEM_BYTES) != cudaSuccess || per_sm < 1) per_sm = 1;
  kern<<<ctx->sm_count * per_sm, THREADS, SMEM_BYTES, ctx->stream>>>(a);
  return R1C_OK;
}

}  // namespace

// put / prep of a device job list through the TMA tile kernel.  *handled = 0: this path cannot take the plane (row
// pitch or base not 16-byte aligned, no driver entry point) and the caller uses mc.cu's first-generation kernel.
int r1c_mc_tma_launch(r1c_ctx *ctx, int kind, r1c_plane *dst, int16_t *d_tmp, const r1c_plane *src, const r1c_mc_job *d_jobs,
                      int64_t n, int bit_depth, int *handled) {
  *handled = 0;
  if (((src->cfg.stride * src->px_bytes) & 15) != 0 || ((uintptr_t)src->d_data & 15) != 0 || n >= (1ll << 26)) return R1C_OK;
  TmaArgs a;
  if (!get_maps(src, ctx->device, a.map)) return R1C_OK;
  *handled = 1;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  // unit list: one entry per job, plus one per extra 32x16 tile of the jobs larger than that (at most 32 per job)
  void *d_counts, *d_units;
  int rc;
  if ((rc = r1c_scratch_dev(ctx, 2, (2 * NCLS + 4) * sizeof(uint32_t), &d_counts))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 3, (size_t)n * 32 * sizeof(uint32_t), &d_units))) return rc;
  R1C_CUDA(ctx, cudaMemsetAsync(d_counts, 0, (2 * NCLS + 4) * sizeof(uint32_t), ctx->stream));
  const unsigned nb = (unsigned)((n + 255) / 256);
  mc_count_kernel<<<nb, 256, 0, ctx->stream>>>(d_jobs, n, (uint32_t *)d_counts);
  R1C_LAUNCH_CHECK(ctx);
  mc_scatter_kernel<<<nb, 256, 0, ctx->stream>>>(d_jobs, n, (uint32_t *)d_counts, (uint32_t *)d_units);
  R1C_LAUNCH_CHECK(ctx);
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
  a.units = (const uint32_t *)d_units;
  a.cls_count = (const uint32_t *)d_counts;
  a.n = n;
  a.bit_depth = bit_depth;
  r1c_prof_begin(ctx, ctx->prof_mc_slot ? ctx->prof_mc_slot : R1C_PROF_MC);
  if (src->px_bytes == 1) {
    if (kind == 0) rc = launch_kernel<1, 0>(ctx, a);
    else if (kind == 1) rc = launch_kernel<1, 1>(ctx, a);
    else rc = launch_kernel<1, 2>(ctx, a);
  } else {
    if (kind == 0) rc = launch_kernel<2, 0>(ctx, a);
    else if (kind == 1) rc = launch_kernel<2, 1>(ctx, a);
    else rc = launch_kernel<2, 2>(ctx, a);
  }
  r1c_prof_end(ctx, ctx->prof_mc_slot ? ctx->prof_mc_slot : R1C_PROF_MC);
  if (rc) return rc;
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}
