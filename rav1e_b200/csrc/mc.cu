// mc.cu -- batched sub-pel motion compensation: put_8tap / prep_8tap / mc_avg
// (reference src/mc.rs:410-706; semantics of put_8tap_ref :213-307,
// prep_8tap_ref :310-387, mc_avg_ref :389-408, get_filter :201-210).
//
// Work decomposition.  A job (one block, 2..128 px on a side) is cut into
// ITEMS of 4 columns x up to 16 rows; one thread owns one item and walks down
// its rows with the 8-row window of horizontally filtered values held in
// registers: per row 4 x 8 horizontal MACs, then 4 x 8 vertical MACs (DP4A /
// DP2A on packed pixels and packed i16 pairs).  Source rows of the vector
// walker arrive through a per-thread ring in shared memory filled by cp.async,
// 8 rows ahead; strips it cannot take (fewer than 4 columns, unaligned output,
// footprint on the edge of the allocation) go through a scalar per-pixel path.
// Item counts per job go through a prefix sum (cub) so that job lists mixing
// every block size still fill all lanes; the main kernel's resident CTAs take
// chunks of items from a device counter and read the item total from device
// memory, so the `_dev` entry points never synchronise with the host.
//
// Arithmetic follows the `_ref` functions literally: i32 accumulation, the
// horizontal intermediate truncated to i16 (`as i16`, src/mc.rs:281,366),
// intermediate_bits = 4 (8/10-bit) or 2 (12-bit), and the four (col_frac,
// row_frac) cases with their own rounding.
// Algorithmic bytes: put/prep 2*px_bytes per output pixel (1 read + 1 write,
// halo served by L1/L2); mc_avg 2*2 B read + px_bytes written.
#include <stdlib.h>

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

constexpr int SEG_ROWS = 16;  // rows per item
constexpr int MC_THREADS = 128;
#ifndef MC_PF
#define MC_PF 8  // source rows in flight per thread in the fast strip walker
#endif

__host__ __device__ inline int items_of(int w, int h) { return ((w + 3) >> 2) * ((h + SEG_ROWS - 1) / SEG_ROWS); }

// get_filter, src/mc.rs:201-210
__device__ __forceinline__ int filter_bank(int mode, int length) {
  return (mode == R1C_BILINEAR || length > 4) ? mode : ((mode < 1 ? mode : 1) + 4);
}

__device__ __forceinline__ int rshift_round(int v, int bit) { return (v + ((1 << bit) >> 1)) >> bit; }

template <typename JOB, int ROWS_PER_ITEM>
__global__ void count_items_kernel(const JOB *jobs, long long n, long long *counts, long long *minmax) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const long long c = ROWS_PER_ITEM == 1 ? (long long)((jobs[i].w + 3) >> 2) * jobs[i].h : items_of(jobs[i].w, jobs[i].h);
    counts[i] = c;
    // min / max items per job: equal -> the kernels index jobs by division instead of searching
    const long long wmin = __reduce_min_sync(__activemask(), (unsigned)c), wmax = __reduce_max_sync(__activemask(), (unsigned)c);
    if ((threadIdx.x & 31) == 0 || i == 0) {
      atomicMin(minmax, wmin);
      atomicMin(minmax + 1, -wmax);  // kept negated so one memset initialises both
    }
  }
  if (i == n) {
    counts[i] = 0;
    minmax[2] = 0;  // work counter of the main kernel (chunks of items handed to CTAs)
  }
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

struct McArgs {
  PlaneView src;
  uint8_t *dst0;  // put: address of dst plane coordinate (0,0)
  int dst_stride;
  int16_t *tmp;   // prep
  const r1c_mc_job *jobs;
  long long *offs;
  long long n;
  int bit_depth;
};

// ---- packed-arithmetic strip walker -------------------------------------------------
// Taps are s8 (every filter row with frac != 0 has |tap| <= 126), so an 8-tap sum is
// 2 DP4A on u8 pixels or 4 DP2A on u16 pixels / i16 intermediates:
//   t0 = taps 0..3, t1 = taps 4..7, one s8 per byte.
__device__ __forceinline__ int dp4a_u8s8(uint32_t px, uint32_t taps, int acc) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_lo_u(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.lo.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_hi_u(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.hi.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_lo_s(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ int dp2a_hi_s(uint32_t pair, uint32_t taps, int acc) {
  int d;
  asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(taps), "r"(acc));
  return d;
}
__device__ __forceinline__ uint32_t pack_taps4(const short *f) {
  return ((uint32_t)(uint8_t)(int8_t)f[0]) | ((uint32_t)(uint8_t)(int8_t)f[1] << 8) |
         ((uint32_t)(uint8_t)(int8_t)f[2] << 16) | ((uint32_t)(uint8_t)(int8_t)f[3] << 24);
}

// raw aligned words of one source row of a strip: the 4 (u8) / 6 (u16) words from the one holding pixel x0-3 on
template <int PB>
struct RowWords {
  uint32_t w[PB == 1 ? 4 : 6];
};
// horizontal 8-tap sums (+ rnd) of the 4 columns of the strip.  `mis` = misalignment of the
// row start inside its first word: bytes (PB 1) or 0 / 1 pixel (PB 2); the same for every
// row of an item (strides are multiples of 4 bytes).
template <int PB>
__device__ __forceinline__ void hsum_words(const RowWords<PB> &rw, uint32_t mis, uint32_t t0, uint32_t t1, int rnd,
                                           int (&s)[4]) {
  if (PB == 1) {
    const uint32_t sh = mis * 8;
    const uint32_t a = __funnelshift_r(rw.w[0], rw.w[1], sh), b = __funnelshift_r(rw.w[1], rw.w[2], sh),
                   c = __funnelshift_r(rw.w[2], rw.w[3], sh);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t lo = k ? __funnelshift_r(a, b, 8 * k) : a, hi = k ? __funnelshift_r(b, c, 8 * k) : b;
      s[k] = dp4a_u8s8(hi, t1, dp4a_u8s8(lo, t0, rnd));
    }
  } else {
    // realign so that pixel x0-3 is the low half of v[0], then q = the odd-phase pairs
    const uint32_t sh = mis * 16;
    uint32_t v[6], q[5];
#pragma unroll
    for (int i = 0; i < 6; i++) v[i] = __funnelshift_r(rw.w[i], i < 5 ? rw.w[i + 1] : 0u, sh);
#pragma unroll
    for (int i = 0; i < 5; i++) q[i] = __byte_perm(v[i], v[i + 1], 0x5432);  // (hi(v[i]), lo(v[i+1]))
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t *pp = (k & 1) ? &q[k >> 1] : &v[k >> 1];
      s[k] = dp2a_hi_u(pp[3], t1, dp2a_lo_u(pp[2], t1, dp2a_hi_u(pp[1], t0, dp2a_lo_u(pp[0], t0, rnd))));
    }
  }
}

// halved taps as s8 x 4 (every SUBPEL_FILTERS entry is even; frac 0 = the unit tap 128 -> 64), packed once per
// launch chain by pack_taps_kernel: lanes of a warp hold different (bank, frac), which serialises constant-bank reads
__device__ uint2 g_half_taps[6 * 16];
__global__ void pack_taps_kernel() {
  const int i = threadIdx.x;
  if (i >= 6 * 16) return;
  const short *f = c_filters[i >> 4][i & 15];
  short h[8];
#pragma unroll
  for (int k = 0; k < 8; k++) h[k] = (short)(f[k] >> 1);
  g_half_taps[i] = make_uint2(pack_taps4(h), pack_taps4(h + 4));
}
__device__ __forceinline__ void half_taps(int bank, int frac, uint32_t &t0, uint32_t &t1) {
  const uint2 t = __ldg(&g_half_taps[bank * 16 + (frac & 15)]);
  t0 = t.x;
  t1 = t.y;
}

// One strip (<= 4 columns) of rows [r0, r1) of a job, pixel by pixel: the kernel's path for strips the vector walker
// below cannot take (fewer than 4 live columns, an unaligned output row, or a source footprint that touches the
// edge of the plane allocation).  ONE formula for the four (col_frac, row_frac) cases of put_8tap_ref /
// prep_8tap_ref (src/mc.rs:224-306, 326-386): a zero frac is the unit filter (tap 128), for which the separable
// formula  mid = round_shift(sum_x, 7-ib) as i16;  out = round_shift(sum_y, 7+ib | 7)  reproduces the reference's
// 1-D and copy cases bit for bit (128 * px >> (7-ib) = px << ib exactly; round_shift(128 * mid, 7+ib) =
// round_shift(mid, ib)).  Taps are halved (all are even), so every shift drops by one:
// round_shift(2 S, s) = (S + 2^(s-2)) >> (s-1).  Pixels under a zero tap are not read, so a case never touches
// rows or columns the reference does not.
template <int PB, int KIND>
__device__ __forceinline__ void walk_strip_scalar(const McArgs &a, const r1c_mc_job &j, int x0, int ncols, int r0, int r1,
                                               int ib, int maxv) {
  const short *fx = c_filters[filter_bank(j.mode_x, j.w)][j.col_frac & 15];
  const short *fy = c_filters[filter_bank(j.mode_y, j.h)][j.row_frac & 15];
  const int hsh = 6 - ib, hrnd = 1 << (5 - ib);
  const int vshift = (KIND == 1 ? 7 : 7 + ib) - 1, vrnd = 1 << (vshift - 1);
  const uint8_t *src = (const uint8_t *)a.src.p0 + ((long long)j.src_y * a.src.stride + j.src_x + x0) * PB;
  const long long sstride = (long long)a.src.stride * PB;
  for (int r = r0; r < r1; r++)
    for (int c = 0; c < ncols; c++) {
      int sv = vrnd;
      for (int ky = 0; ky < 8; ky++) {
        const int ty = fy[ky] >> 1;
        if (ty == 0) continue;
        const uint8_t *row = src + (long long)(r + ky - 3) * sstride;
        int sh = hrnd;
        for (int kx = 0; kx < 8; kx++) {
          const int tx = fx[kx] >> 1;
          if (tx == 0) continue;
          sh += tx * (PB == 1 ? (int)row[c + kx - 3] : (int)((const uint16_t *)row)[c + kx - 3]);
        }
        sv += ty * (int)(int16_t)(sh >> hsh);
      }
      const int v = sv >> vshift;
      if (KIND == 1) {
        a.tmp[(long long)j.tmp_off + (long long)r * j.w + x0 + c] = (int16_t)v;
      } else {
        const int o = min(max(v, 0), maxv);
        uint8_t *d = KIND == 0 ? a.dst0 + ((long long)(j.dst_y + r) * a.dst_stride + j.dst_x + x0 + c) * PB
                               : (uint8_t *)a.tmp + ((long long)j.tmp_off + (long long)r * j.w + x0 + c) * PB;
        if (PB == 1) *d = (uint8_t)o;
        else *(uint16_t *)d = (uint16_t)o;
      }
    }
}

// The same strip for the common case -- 4 live columns, an output row that can be stored as one aligned vector, and
// a source footprint (all 6 / 4 aligned words of rows r0-3 .. r1+3) that lies inside the plane allocation whatever
// the (col_frac, row_frac) case: words and rows a case does not need are read anyway and meet zero taps.  No
// per-word predicates, row pointers advance by the stride, outputs are clamped and stored as packed pairs.
template <int PB, int KIND, int PF>
__device__ __forceinline__ void walk_strip_fast(const McArgs &a, const r1c_mc_job &j, int x0, int r0, int r1, int ib,
                                                int maxv) {
  uint32_t tx0, tx1, ty0, ty1;
  half_taps(filter_bank(j.mode_x, j.w), j.col_frac, tx0, tx1);
  half_taps(filter_bank(j.mode_y, j.h), j.row_frac, ty0, ty1);
  const int hsh = 6 - ib, hrnd = 1 << (5 - ib);
  const int vshift = (KIND == 1 ? 7 : 7 + ib) - 1, vrnd = 1 << (vshift - 1);
  const long long sstride = (long long)a.src.stride * PB;
  const uint8_t *hfirst = (const uint8_t *)a.src.p0 + ((long long)(j.src_y + r0 - 3) * a.src.stride + j.src_x + x0 - 3) * PB;
  const uint32_t mis = PB == 1 ? (uint32_t)((uintptr_t)hfirst & 3) : (uint32_t)(((uintptr_t)hfirst >> 1) & 1);
  const uint8_t *wp = (const uint8_t *)((uintptr_t)hfirst & ~(uintptr_t)3);  // next row to fetch
  constexpr int NW = PB == 1 ? 4 : 6;
  const int nrows = r1 - r0 + 7;
  uint8_t *dp;  // output row
  long long dstride;
  if (KIND == 0) {
    dp = a.dst0 + ((long long)(j.dst_y + r0) * a.dst_stride + j.dst_x + x0) * PB;
    dstride = (long long)a.dst_stride * PB;
  } else {
    const int ob = KIND == 1 ? 2 : PB;
    dp = (uint8_t *)a.tmp + ((long long)j.tmp_off + (long long)r0 * j.w + x0) * ob;
    dstride = (long long)j.w * ob;
  }
  // Source rows travel through a per-thread ring in shared memory filled by cp.async, PF rows ahead, one commit group
  // per row.  (A ring of registers filled by plain loads does not work: a warp has six scoreboards, the loads of
  // several rows share one, and the wait for the oldest row then also waits for the youngest -- ncu showed every
  // row paying a full memory latency.)  wait_group<PF - 1> is exact: PF + i groups exist when row i is needed.
  __shared__ uint32_t s_ring[PF * NW * MC_THREADS];
  uint32_t *mine = s_ring + threadIdx.x;
  auto fetch = [&](int slot) {
#pragma unroll
    for (int k = 0; k < NW; k++) {
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(mine + (slot * NW + k) * MC_THREADS);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(wp + 4 * k) : "memory");
    }
    wp += sstride;
  };
#pragma unroll
  for (int k = 0; k < PF; k++) {
    if (k < nrows) fetch(k);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  uint32_t pr[4][8];
  int prev[4] = {0, 0, 0, 0};
  const uint32_t mx = (uint32_t)maxv * 0x10001u;
  for (int ib8 = 0; ib8 < nrows; ib8 += 8) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = ib8 + u;  // source row r0 - 3 + i
      if (i < nrows) {
        asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");
        RowWords<PB> cur;
#pragma unroll
        for (int k = 0; k < NW; k++) cur.w[k] = mine[((u % PF) * NW + k) * MC_THREADS];
        int sum[4];
        hsum_words<PB>(cur, mis, tx0, tx1, hrnd, sum);
        if (i + PF < nrows) fetch(u % PF);
        asm volatile("cp.async.commit_group;" ::: "memory");
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const int mid = sum[c] >> hsh;
          pr[c][(u + 7) & 7] = __byte_perm((uint32_t)prev[c], (uint32_t)mid, 0x5410);
          prev[c] = mid;
        }
        if (i >= 7) {
          int v[4];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            int t = dp2a_lo_s(pr[c][(u + 1) & 7], ty0, vrnd);
            t = dp2a_hi_s(pr[c][(u + 3) & 7], ty0, t);
            t = dp2a_lo_s(pr[c][(u + 5) & 7], ty1, t);
            t = dp2a_hi_s(pr[c][(u + 7) & 7], ty1, t);
            v[c] = t >> vshift;
          }
          // every result fits i16 (a pixel plus filter overshoot; prep keeps the low 16 bits like `as i16`)
          uint32_t p01 = __byte_perm((uint32_t)v[0], (uint32_t)v[1], 0x5410), p23 = __byte_perm((uint32_t)v[2], (uint32_t)v[3], 0x5410);
          if (KIND != 1) {
            p01 = __vmins2(__vmaxs2(p01, 0u), mx);
            p23 = __vmins2(__vmaxs2(p23, 0u), mx);
          }
          if (KIND != 1 && PB == 1) *(uint32_t *)dp = __byte_perm(p01, p23, 0x6420);
          else *(uint2 *)dp = make_uint2(p01, p23);
          dp += dstride;
        }
      }
    }
  }
}

// KIND 0 = put (clamped pixels into the dst plane), 1 = prep (i16, dense w*h at tmp_off),
// 2 = put into a dense w*h pixel block at tmp_off (pixels; used by the sub-pel search)
template <int PB, int KIND>
__global__ void __launch_bounds__(MC_THREADS, 4) mc_kernel(const __grid_constant__ McArgs a) {
  const long long total = a.offs[a.n];
  const int ipj = a.offs[a.n + 1] == -a.offs[a.n + 2] ? (int)a.offs[a.n + 1] : 0;  // min == max items per job
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;
  // resident CTAs take chunks of MC_THREADS items from a counter: strips differ a lot in cost (rows, fast / scalar
  // path), and a static split left SMs idle for the last third of the launch
  __shared__ long long s_base;
  unsigned long long *counter = (unsigned long long *)(a.offs + a.n + 3);
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_base = (long long)atomicAdd(counter, (unsigned long long)MC_THREADS);
    __syncthreads();
    const long long item = s_base + threadIdx.x;
    if (s_base >= total) break;
    if (item >= total) continue;
    // every job with the same number of items (the usual batch): no search
    const long long ji = ipj ? (long long)((unsigned long long)item / (unsigned)ipj) : find_job(a.offs, a.n, item);
    const r1c_mc_job j = a.jobs[ji];
    const int local = ipj ? (int)(item - ji * ipj) : (int)(item - a.offs[ji]);
    const int nstrips = (j.w + 3) >> 2;
    const int seg = local / nstrips, strip = local - seg * nstrips;
    const int x0 = strip * 4;
    const int ncols = min(4, j.w - x0);
    const int r0 = seg * SEG_ROWS, r1 = min((int)j.h, r0 + SEG_ROWS);
    // footprint [x0 - 4, x0 + 12) x [r0 - 3, r1 + 3] of the source inside the allocation, vector store aligned
    const int fx = j.src_x + x0, fy = j.src_y;
    const bool src_safe = fx - 4 >= -a.src.xorigin && fx + 12 <= a.src.alloc_w - a.src.xorigin &&
                          fy + r0 - 3 >= -a.src.yorigin && fy + r1 + 3 < a.src.alloc_h - a.src.yorigin;
    const int ob = KIND == 1 ? 2 : PB;  // bytes per output element
    const uintptr_t daddr = KIND == 0 ? (uintptr_t)a.dst0 + (uintptr_t)(((long long)j.dst_y * a.dst_stride + j.dst_x + x0) * PB)
                                      : (uintptr_t)a.tmp + (uintptr_t)(((long long)j.tmp_off + x0) * ob);
    const int drow = KIND == 0 ? a.dst_stride * PB : j.w * ob;
    if (ncols == 4 && src_safe && ((daddr | (uintptr_t)drow) & (uintptr_t)(4 * ob - 1)) == 0)
      walk_strip_fast<PB, KIND, MC_PF>(a, j, x0, r0, r1, ib, maxv);
    else
      walk_strip_scalar<PB, KIND>(a, j, x0, ncols, r0, r1, ib, maxv);
  }
}

struct AvgArgs {
  uint8_t *dst0;
  int dst_stride;
  const int16_t *t1, *t2;
  const r1c_avg_job *jobs;
  long long n;
  int bit_depth;
  int lg_g;  // a warp takes 1 << lg_g consecutive jobs
};

// what a lane needs to find the job of an item of its warp's chunk
struct AvgChunk {
  const uint4 *job;  // shared memory: the chunk's jobs
  const int *pre;    // shared memory: exclusive prefix sums of their item counts (INT_MAX behind the last job)
  int total, c0, lg_c0;
  bool uniform;
};

// The items of a chunk, 1 << LGP pixels of one row each, U per lane and trip: all their loads are issued before the
// arithmetic (the kernel lasts only a few DRAM latencies, so the number of independent requests in flight per thread
// is what sets its bandwidth).  LGP = 2: 8-byte loads / stores, U = 4; LGP = 3 (every width of the chunk a multiple of
// 8): 16-byte ones, U = 2 -- the same bytes in flight and registers, half the index arithmetic per pixel (ncu: the
// 4-pixel walker alone kept the ALU pipe 68 % busy at 0.62 of the HBM peak).
template <int PB, int LGP, int U>
__device__ __forceinline__ void avg_walk(const AvgArgs &a, const AvgChunk &ck, int lane) {
  constexpr int NP = 1 << LGP, NWD = NP / 2;  // pixels / 32-bit words of intermediates per item
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;
  const uintptr_t tbase = ((uintptr_t)a.t1) | ((uintptr_t)a.t2);
  const bool base_ok = (tbase & (2 * NP - 1)) == 0;  // vector loads need both arrays aligned
  for (int item0 = lane; item0 < ck.total; item0 += 32 * U) {
    long long o[U];
    uint8_t *d[U];
    int ncols[U];
    bool live[U], vec[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int item = item0 + 32 * u;
      live[u] = item < ck.total;
      const int it = live[u] ? item : item0;
      int ji;
      if (ck.lg_c0 >= 0) ji = it >> ck.lg_c0;
      else if (ck.uniform) ji = (int)((unsigned)it / (unsigned)ck.c0);
      else {
        ji = 0;  // largest j with pre[j] <= it
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
          if (ck.pre[ji + s] <= it) ji += s;
      }
      const uint4 jb = ck.job[ji];  // dst_x, dst_y, w | h << 16, tmp_off
      const long long toff = (long long)(int)jb.w;
      const int jx = (int)jb.x, jy = (int)jb.y, w = (int)(int16_t)(jb.z & 0xffffu);
      const int local = it - (ck.uniform ? ji * ck.c0 : ck.pre[ji]);
      const int nstrips = (w + NP - 1) >> LGP;
      const int r = (nstrips & (nstrips - 1)) == 0 ? local >> (__ffs(nstrips) - 1) : local / nstrips;
      const int strip = local - r * nstrips;
      const int x0 = strip << LGP;
      ncols[u] = min(NP, w - x0);
      o[u] = toff + (long long)r * w + x0;
      d[u] = a.dst0 + ((long long)(jy + r) * a.dst_stride + jx + x0) * PB;
      vec[u] = base_ok && ncols[u] == NP && (o[u] & (NP - 1)) == 0;
    }
    uint32_t p[U][NWD], q[U][NWD];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (live[u] && vec[u]) {
        if (LGP == 3) {
          const uint4 pv = __ldg((const uint4 *)(a.t1 + o[u])), qv = __ldg((const uint4 *)(a.t2 + o[u]));
          p[u][0] = pv.x, p[u][1] = pv.y, p[u][NWD - 2] = pv.z, p[u][NWD - 1] = pv.w;
          q[u][0] = qv.x, q[u][1] = qv.y, q[u][NWD - 2] = qv.z, q[u][NWD - 1] = qv.w;
        } else {
          const uint2 pv = __ldg((const uint2 *)(a.t1 + o[u])), qv = __ldg((const uint2 *)(a.t2 + o[u]));
          p[u][0] = pv.x, p[u][1] = pv.y;
          q[u][0] = qv.x, q[u][1] = qv.y;
        }
      }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (!live[u]) continue;
      int v[NP];
      if (vec[u]) {
#pragma unroll
        for (int k = 0; k < NWD; k++) {
          v[2 * k] = (int)(int16_t)p[u][k] + (int)(int16_t)q[u][k];
          v[2 * k + 1] = ((int)p[u][k] >> 16) + ((int)q[u][k] >> 16);
        }
      } else {
#pragma unroll
        for (int c = 0; c < NP; c++) v[c] = c < ncols[u] ? (int)a.t1[o[u] + c] + (int)a.t2[o[u] + c] : 0;
      }
#pragma unroll
      for (int c = 0; c < NP; c++) v[c] = min(max(rshift_round(v[c], ib + 1), 0), maxv);
      uint8_t *dp = d[u];
      if (ncols[u] == NP && ((uintptr_t)dp & (NP * PB - 1)) == 0) {
        uint32_t wd[NP * PB / 4];
#pragma unroll
        for (int k = 0; k < NP * PB / 4; k++)
          wd[k] = PB == 1 ? ((uint32_t)v[4 * k] | ((uint32_t)v[4 * k + 1] << 8) | ((uint32_t)v[4 * k + 2] << 16) | ((uint32_t)v[4 * k + 3] << 24))
                          : ((uint32_t)v[2 * k] | ((uint32_t)v[2 * k + 1] << 16));
        if (NP * PB == 4) *(uint32_t *)dp = wd[0];
        else if (NP * PB == 8) *(uint2 *)dp = make_uint2(wd[0], wd[1]);
        else *(uint4 *)dp = make_uint4(wd[0], wd[1], wd[(NP * PB / 4) - 2], wd[(NP * PB / 4) - 1]);
      } else {
#pragma unroll
        for (int c = 0; c < NP; c++)
          if (c < ncols[u]) {
            if (PB == 1) dp[c] = (uint8_t)v[c];
            else ((uint16_t *)dp)[c] = (uint16_t)v[c];
          }
      }
    }
  }
}

// The same for a chunk whose jobs are all 4 pixels wide with an even height (4x4, 4x8, 4x16): an item is TWO rows of a
// block = 8 consecutive intermediates (one 16-byte load per array) and two 4-pixel stores a row apart.
template <int PB, int U>
__device__ __forceinline__ void avg_walk_w4(const AvgArgs &a, const AvgChunk &ck, int lane) {
  const int ib = a.bit_depth == 12 ? 2 : 4;
  const int maxv = (1 << a.bit_depth) - 1;
  const bool base_ok = ((((uintptr_t)a.t1) | ((uintptr_t)a.t2)) & 15) == 0;
  const long long row_bytes = (long long)a.dst_stride * PB;
  for (int item0 = lane; item0 < ck.total; item0 += 32 * U) {
    long long o[U];
    uint8_t *d[U];
    bool live[U], vec[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int item = item0 + 32 * u;
      live[u] = item < ck.total;
      const int it = live[u] ? item : item0;
      int ji;
      if (ck.lg_c0 >= 0) ji = it >> ck.lg_c0;
      else if (ck.uniform) ji = (int)((unsigned)it / (unsigned)ck.c0);
      else {
        ji = 0;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
          if (ck.pre[ji + s] <= it) ji += s;
      }
      const uint4 jb = ck.job[ji];
      const int local = it - (ck.uniform ? ji * ck.c0 : ck.pre[ji]);  // row pair of the block
      o[u] = (long long)(int)jb.w + 8 * local;
      d[u] = a.dst0 + ((long long)((int)jb.y + 2 * local) * a.dst_stride + (int)jb.x) * PB;
      vec[u] = base_ok && (o[u] & 7) == 0;
    }
    uint32_t p[U][4], q[U][4];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (live[u] && vec[u]) {
        const uint4 pv = __ldg((const uint4 *)(a.t1 + o[u])), qv = __ldg((const uint4 *)(a.t2 + o[u]));
        p[u][0] = pv.x, p[u][1] = pv.y, p[u][2] = pv.z, p[u][3] = pv.w;
        q[u][0] = qv.x, q[u][1] = qv.y, q[u][2] = qv.z, q[u][3] = qv.w;
      }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (!live[u]) continue;
      int v[8];
      if (vec[u]) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
          v[2 * k] = (int)(int16_t)p[u][k] + (int)(int16_t)q[u][k];
          v[2 * k + 1] = ((int)p[u][k] >> 16) + ((int)q[u][k] >> 16);
        }
      } else {
#pragma unroll
        for (int c = 0; c < 8; c++) v[c] = (int)a.t1[o[u] + c] + (int)a.t2[o[u] + c];
      }
#pragma unroll
      for (int c = 0; c < 8; c++) v[c] = min(max(rshift_round(v[c], ib + 1), 0), maxv);
#pragma unroll
      for (int rr = 0; rr < 2; rr++) {
        uint8_t *dp = d[u] + rr * row_bytes;
        const int *vr = v + 4 * rr;
        if (((uintptr_t)dp & (4 * PB - 1)) == 0) {
          if (PB == 1) *(uint32_t *)dp = (uint32_t)vr[0] | ((uint32_t)vr[1] << 8) | ((uint32_t)vr[2] << 16) | ((uint32_t)vr[3] << 24);
          else *(uint2 *)dp = make_uint2((uint32_t)vr[0] | ((uint32_t)vr[1] << 16), (uint32_t)vr[2] | ((uint32_t)vr[3] << 16));
        } else {
#pragma unroll
          for (int c = 0; c < 4; c++) {
            if (PB == 1) dp[c] = (uint8_t)vr[c];
            else ((uint16_t *)dp)[c] = (uint16_t)vr[c];
          }
        }
      }
    }
  }
}

// mc_avg_ref, src/mc.rs:389-408
// One launch, no scratch: a warp takes a CHUNK of up to 32 consecutive jobs (one per lane), prefix-sums their item
// counts with shuffles and then walks the chunk's items with all its lanes (avg_walk).  (Round 1 counted the items of
// the whole list and ran two device-wide scans before a persistent kernel: for 4x4 blocks that scheduling traffic --
// 40 bytes per job -- was larger than the 96 bytes of pixels the job moves, and the three extra launches were a third
// of the call.)  The host picks the chunk size from n alone so that the grid has at least ~32 warps per SM; the
// hardware block scheduler does the load balancing.
template <int PB>
__global__ void __launch_bounds__(MC_THREADS, 9) avg_kernel(AvgArgs a) {
  constexpr int WARPS = MC_THREADS / 32;
  constexpr unsigned FULL = 0xffffffffu;
  __shared__ uint4 s_job[WARPS][32];
  __shared__ int s_pre[WARPS][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long j0 = ((long long)blockIdx.x * WARPS + warp) << a.lg_g;
  if (j0 >= a.n) return;
  const int g = (int)min(1ll << a.lg_g, a.n - j0);
  static_assert(sizeof(r1c_avg_job) == 16, "r1c_avg_job is read as one 16-byte word");
  uint4 jw = make_uint4(0, 0, 8, 0);  // dst_x, dst_y, w | h << 16, tmp_off (padding lanes: w = 8, h = 0)
  if (lane < g) jw = __ldg((const uint4 *)(a.jobs + j0 + lane));
  const int w = (int)(int16_t)(jw.z & 0xffffu), h = (int)(int16_t)(jw.z >> 16);
  // 8 pixels per item when every width of the chunk is a multiple of 8 (one row) or 4 with an even height (two rows),
  // else 4
  const bool wide = __all_sync(FULL, (w & 7) == 0);
  const bool pairs = __all_sync(FULL, lane >= g || (w == 4 && (h & 1) == 0));
  const int c = lane < g ? (wide ? (w >> 3) * h : (pairs ? (h >> 1) : ((w + 3) >> 2) * h)) : 0;
  int inc = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(FULL, inc, d);
    if (lane >= d) inc += t;
  }
  AvgChunk ck;
  ck.total = __shfl_sync(FULL, inc, 31);
  ck.c0 = __shfl_sync(FULL, c, 0);
  // every job of the chunk with the same number of items (the usual batch): a shift or a division, no search
  ck.uniform = __all_sync(FULL, lane >= g || c == ck.c0) && ck.c0 > 0;
  ck.lg_c0 = ck.uniform && (ck.c0 & (ck.c0 - 1)) == 0 ? __ffs(ck.c0) - 1 : -1;
  s_job[warp][lane] = jw;
  s_pre[warp][lane] = lane < g ? inc - c : 0x7fffffff;
  __syncwarp();
  ck.job = s_job[warp];
  ck.pre = s_pre[warp];
  if (wide) avg_walk<PB, 3, 2>(a, ck, lane);  // (3 or 4 items per trip: 72 / 87 registers, slower everywhere)
  else if (pairs) avg_walk_w4<PB, 2>(a, ck, lane);
  else avg_walk<PB, 2, 4>(a, ck, lane);
}

template <typename JOB, int ROWS_PER_ITEM>
int build_offsets(r1c_ctx *ctx, const JOB *d_jobs, int64_t n, long long **d_offs_out) {
  void *d_counts, *d_offs, *d_tmp;
  int rc;
  const size_t bytes = (size_t)(n + 4) * sizeof(long long);  // offs[n+1], offs[n+2] = min / max items per job, offs[n+3] = work counter
  if ((rc = r1c_scratch_dev(ctx, 2, bytes, &d_counts))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 3, bytes, &d_offs))) return rc;
  long long *minmax = (long long *)d_offs + n + 1;
  R1C_CUDA(ctx, cudaMemsetAsync(minmax, 0x7f, 2 * sizeof(long long), ctx->stream));
  count_items_kernel<JOB, ROWS_PER_ITEM><<<(unsigned)((n + 1 + 255) / 256), 256, 0, ctx->stream>>>(d_jobs, n, (long long *)d_counts, minmax);
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

// persistent grids: exactly the CTAs that are resident at once
template <typename K>
int grid_for(r1c_ctx *ctx, K kern) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MC_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
  return ctx->sm_count * per_sm;
}

int launch_mc(r1c_ctx *ctx, int kind, r1c_plane *dst, int16_t *d_tmp, const r1c_plane *src,
              const r1c_mc_job *d_jobs, int64_t n, int bit_depth) {
  if (n == 0) return R1C_OK;
  // the u16 horizontal filter pairs pixels on 32-bit words: rows must keep the parity
  R1C_REQUIRE(ctx, ((src->cfg.stride * src->px_bytes) & 3) == 0, R1C_ERR_ARG, "plane rows must be a multiple of 4 bytes");
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!ctx->mc_taps_ready) {  // identical values whichever context writes them
    pack_taps_kernel<<<1, 128, 0, ctx->stream>>>();
    R1C_LAUNCH_CHECK(ctx);
    ctx->mc_taps_ready = 1;
  }
  {  // second generation: source windows staged by 2-D TMA (mc_tma.cu); R1C_MC_V1=1 keeps this file's kernel
    const char *v1 = getenv("R1C_MC_V1");
    if (!(v1 && v1[0] == '1')) {
      int handled = 0;
      const int rc2 = r1c_mc_tma_launch(ctx, kind, dst, d_tmp, src, d_jobs, n, bit_depth, &handled);
      if (rc2 || handled) return rc2;
    }
  }
  long long *d_offs;
  int rc = build_offsets<r1c_mc_job, 0>(ctx, d_jobs, n, &d_offs);
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
  r1c_prof_begin(ctx, ctx->prof_mc_slot ? ctx->prof_mc_slot : R1C_PROF_MC);
  if (src->px_bytes == 1) {
    if (kind == 0) mc_kernel<1, 0><<<grid_for(ctx, mc_kernel<1, 0>), MC_THREADS, 0, ctx->stream>>>(a);
    else if (kind == 1) mc_kernel<1, 1><<<grid_for(ctx, mc_kernel<1, 1>), MC_THREADS, 0, ctx->stream>>>(a);
    else mc_kernel<1, 2><<<grid_for(ctx, mc_kernel<1, 2>), MC_THREADS, 0, ctx->stream>>>(a);
  } else {
    if (kind == 0) mc_kernel<2, 0><<<grid_for(ctx, mc_kernel<2, 0>), MC_THREADS, 0, ctx->stream>>>(a);
    else if (kind == 1) mc_kernel<2, 1><<<grid_for(ctx, mc_kernel<2, 1>), MC_THREADS, 0, ctx->stream>>>(a);
    else mc_kernel<2, 2><<<grid_for(ctx, mc_kernel<2, 2>), MC_THREADS, 0, ctx->stream>>>(a);
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
  const void *h_jobs;
  void *d_jobs;
  const size_t jb = (size_t)n * sizeof(r1c_mc_job);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_host_src(ctx, 0, jobs, jb, &h_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
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
  const void *h_jobs, *h_in;
  void *d_jobs, *h_tmp, *d_tmp;
  int staged;
  const size_t jb = (size_t)n * sizeof(r1c_mc_job), tb = (size_t)tmp_len * 2;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  // page-locked caller buffers are the DMA source / target themselves; elements no job writes keep the caller's values
  if ((rc = r1c_host_src(ctx, 0, jobs, jb, &h_jobs))) return rc;
  if ((rc = r1c_host_src(ctx, 1, tmp, tb, &h_in))) return rc;
  if ((rc = r1c_host_dst(ctx, 1, tmp, tb, &h_tmp, &staged))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, tb, &d_tmp))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_tmp, h_in, tb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_mc(ctx, 1, nullptr, (int16_t *)d_tmp, src, (const r1c_mc_job *)d_jobs, n, bit_depth))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_tmp, d_tmp, tb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (staged) memcpy(tmp, h_tmp, tb);
  return R1C_OK;
}

static int launch_avg(r1c_ctx *ctx, r1c_plane *dst, const int16_t *d_t1, const int16_t *d_t2,
                      const r1c_avg_job *d_jobs, int64_t n, int bit_depth) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  R1C_REQUIRE(ctx, ((uintptr_t)d_jobs & 15) == 0, R1C_ERR_ARG, "mc_avg: the job array must be 16-byte aligned");
  AvgArgs a;
  PlaneView dv = make_view(dst);
  a.dst0 = (uint8_t *)dv.p0;
  a.dst_stride = dv.stride;
  a.t1 = d_t1;
  a.t2 = d_t2;
  a.jobs = d_jobs;
  a.n = n;
  a.bit_depth = bit_depth;
  // jobs per warp: as many as leave ~32 warps per SM (a frame of 64x64 blocks is only ~2000 jobs: one per warp)
  a.lg_g = 5;
  while (a.lg_g > 0 && (n >> a.lg_g) < (int64_t)ctx->sm_count * 32) a.lg_g--;
  const int64_t chunks = (n + (1ll << a.lg_g) - 1) >> a.lg_g;
  const unsigned grid = (unsigned)((chunks + MC_THREADS / 32 - 1) / (MC_THREADS / 32));
  r1c_prof_begin(ctx, R1C_PROF_MC);
  if (dst->px_bytes == 1) avg_kernel<1><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
  else avg_kernel<2><<<grid, MC_THREADS, 0, ctx->stream>>>(a);
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
