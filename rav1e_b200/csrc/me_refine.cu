// me_refine.cu -- the tail of MotionEstimation::motion_estimation (src/me.rs:232-271)
// for every block of a frame, fused into ONE kernel (8-bit planes; 10-bit u16 planes with PB = 2):
//   * with use_satd: lowest_cost := SATD cost of the full-pel winner (get_mv_rd_cost
//     with tmp_plane_opt = None, src/me.rs:235-252, 828-836);
//   * telescopic_subpel_search (src/me.rs:858-941): steps 8, 4, 2 (and 1 with
//     high-precision MVs), 8 neighbours per step in (i, j) raster order, strict `<`;
//     each candidate = predict_inter (src/predict.rs:255-338: MV -> offset + 1/16
//     fracs, clamped origin, put_8tap REGULAR/REGULAR, src/mc.rs:213-307) into a
//     block-sized buffer, then SAD | SATD (src/dist.rs) + rate (src/me.rs:840-856).
//
// One CTA = half of the blocks of one 64x64 superblock (all 19 sizes, 849 blocks
// per superblock).  The CTA stages the superblock of `org` and the reference
// window that every candidate of every block can touch (+-(range + 2) px + filter
// taps) in shared memory ONCE and runs all steps without leaving the SM:
//   work unit = (block, candidate, S x S tile), S = the SATD tile edge (4 if the
//   block has a side of 4, else 8); S threads own one tile, one column each.  A thread
//   walks down its column: horizontal 8-tap = 2 DP4A on 8 window bytes (taps are s8),
//   vertical 8-tap = IMADs scattered into the S output accumulators, then the column
//   of differences goes through the in-register vertical Hadamard, a shuffle
//   butterfly across the S threads, |.| and a group reduction; one shared-memory
//   atomicAdd per tile accumulates the block's distortion.  After a barrier one
//   thread per block applies normalisation, rate and the strict-`<` pick in
//   candidate order.
// Nothing goes through HBM between the steps (the previous composition of
// kernels wrote every predicted block to a dense buffer: ~2 GB per 1080p frame).
#include <stdlib.h>

#include "dist_dev.cuh"
#include "me_common.cuh"

namespace {

// threads per CTA: 2 CTAs of 256 per SM for u8 planes; the u16 window (83 KB for +-64) leaves room for one CTA of 512
template <int PB>
struct Rc {
  static constexpr int RT = PB == 1 ? 256 : 512;
};
constexpr int NSLOT = 849;     // blocks per superblock
constexpr int SPLIT = 2;       // CTAs per superblock
constexpr int MARGIN = 2;      // sub-pel steps move the MV by < 2 px in total

// REGULAR 8-tap (bank 0) and its 4-tap variant for lengths <= 4 (bank 4),
// SUBPEL_FILTERS of src/mc.rs:70-179, [bank][frac][tap]
__constant__ signed char c_taps[2][16][8] = {
    {{0, 0, 0, 127, 0, 0, 0, 0}, {0, 2, -6, 126, 8, -2, 0, 0}, {0, 2, -10, 122, 18, -4, 0, 0}, {0, 2, -12, 116, 28, -8, 2, 0},
     {0, 2, -14, 110, 38, -10, 2, 0}, {0, 2, -14, 102, 48, -12, 2, 0}, {0, 2, -16, 94, 58, -12, 2, 0}, {0, 2, -14, 84, 66, -12, 2, 0},
     {0, 2, -14, 76, 76, -14, 2, 0}, {0, 2, -12, 66, 84, -14, 2, 0}, {0, 2, -12, 58, 94, -16, 2, 0}, {0, 2, -12, 48, 102, -14, 2, 0},
     {0, 2, -10, 38, 110, -14, 2, 0}, {0, 2, -8, 28, 116, -12, 2, 0}, {0, 0, -4, 18, 122, -10, 2, 0}, {0, 0, -2, 8, 126, -6, 2, 0}},
    {{0, 0, 0, 127, 0, 0, 0, 0}, {0, 0, -4, 126, 8, -2, 0, 0}, {0, 0, -8, 122, 18, -4, 0, 0}, {0, 0, -10, 116, 28, -6, 0, 0},
     {0, 0, -12, 110, 38, -8, 0, 0}, {0, 0, -12, 102, 48, -10, 0, 0}, {0, 0, -14, 94, 58, -10, 0, 0}, {0, 0, -12, 84, 66, -10, 0, 0},
     {0, 0, -12, 76, 76, -12, 0, 0}, {0, 0, -10, 66, 84, -12, 0, 0}, {0, 0, -10, 58, 94, -14, 0, 0}, {0, 0, -10, 48, 102, -12, 0, 0},
     {0, 0, -8, 38, 110, -12, 0, 0}, {0, 0, -6, 28, 116, -10, 0, 0}, {0, 0, -4, 18, 122, -8, 0, 0}, {0, 0, -2, 8, 126, -4, 0, 0}}};
// (row 0 is a placeholder: the unit tap 128 does not fit s8; the kernel keeps HALVED taps in shared
//  memory, where frac 0 is the unit tap 64 -- the reference copies instead, src/mc.rs:224-231.)

// tile list of a full superblock, the same for every superblock: one u32 per
// S x S tile = slot | bsize << 10 | bx4 << 15 | by4 << 19 | tx4 << 23 | ty4 << 27
// (block origin and tile offset inside the block in units of 4 px).
struct TileTable {
  int n4, n8;            // tiles of the 4x4 class / of the 8x8 class
  int split4[SPLIT + 1]; // CTA c takes tiles [split[c], split[c+1]) -- block boundaries
  int split8[SPLIT + 1];
  int slot_split[SPLIT + 1][2];
};
constexpr int MAX_T4 = 1280, MAX_T8 = 896;
__device__ uint32_t g_tiles4[MAX_T4];
__device__ uint32_t g_tiles8[MAX_T8];

struct RefArgs {
  PlaneView org, ref;
  const r1c_sb_job *sbs;       // superblock origins (po_x, po_y)
  r1c_me_result *results;      // frame layout, in / out
  long long off[R1C_BLOCK_SIZES_ALL];
  int cols[R1C_BLOCK_SIZES_ALL], rows[R1C_BLOCK_SIZES_ALL];
  TileTable tt;
  int W, H;          // frame size
  int range_x, range_y;  // full-pel search range the MVs come from (window = range + MARGIN)
  uint32_t lambda;
  int use_satd, allow_hp, bit_depth;
  int winW, winH, wstride;  // staged window (bytes per row = wstride, multiple of 4)
};

__host__ __device__ inline int slot_off(int b) {
  int o = 0;
  for (int i = 0; i < b; i++) {
    const int w = blk_w_of(i), h = blk_h_of(i);
    o += (w > 64 || h > 64) ? 0 : (64 / w) * (64 / h);
  }
  return o;
}

struct Blk {       // per-block state in shared memory
  r1c_mv mv;       // best_mv
  r1c_mv center;   // centre of the current step
  unsigned long long cost;
};

__device__ __forceinline__ int rshift_round(int v, int bit) { return (v + ((1 << bit) >> 1)) >> bit; }

__device__ __forceinline__ int dp4a_us(uint32_t px, uint32_t taps, int acc) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(taps), "r"(acc));
  return d;
}

// horizontal 8-tap sum (halved taps, + the rounding `rnd` of the shift that follows: 2 for >> 2, 8 for >> 4 at 12 bits) at window pixel `b` of the row
// starting at `row` (pixels b .. b+7): u8 = 2 DP4A on funnel-shifted words, u16 = 4 DP2A on pixel pairs
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
template <int PB>
__device__ __forceinline__ int hsum8(const uint8_t *row, int b, uint32_t tlo, uint32_t thi, int rnd) {
  if (PB == 1) {
    const uint32_t *w = (const uint32_t *)row + (b >> 2);
    const uint32_t sh = (uint32_t)(b & 3) * 8;
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
    const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
    return dp4a_us(hi, thi, dp4a_us(lo, tlo, rnd));
  } else {
    const uint32_t *w = (const uint32_t *)row + (b >> 1);
    const uint32_t sh = (uint32_t)(b & 1) * 16;
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
    const uint32_t v0 = __funnelshift_r(w0, w1, sh), v1 = __funnelshift_r(w1, w2, sh),
                   v2 = __funnelshift_r(w2, w3, sh), v3 = __funnelshift_r(w3, w4, sh);
    return dp2a_hi_us(v3, thi, dp2a_lo_us(v2, thi, dp2a_hi_us(v1, tlo, dp2a_lo_us(v0, tlo, rnd))));
  }
}

// put_8tap_ref for bit depths 8 and 10 (intermediate_bits = 4, src/mc.rs:213-307) in ONE formula for its four
// (col_frac, row_frac) cases.  Every REGULAR tap is even, so the filters are kept halved (s8) and frac 0
// is the unit tap 64:
//   mid = (sum px * htx + 2) >> 2            == round_shift(sum px * tx, 3), and px << 4 for col_frac 0
//   out = clamp((sum mid * hty + 512) >> 10)  == round_shift(sum mid * ty, 11); round_shift(mid, 4) for
//                                               row_frac 0; round_shift(sum px * ty, 7) for col_frac 0
// which are exactly the reference's four branches (no branch here: the lanes of a warp hold different MVs).
// At 12 bits intermediate_bits = 2: mid = (sum px * htx + 8) >> 4 (px << 2 for col_frac 0) and out = (sum mid * hty + 128)
// >> 8, which collapse onto the reference's branches the same way ((4 A + 256) >> 9 == (A + 64) >> 7; (128 m + 256) >> 9 ==
// (m + 2) >> 2).
// The outer taps (0 and 7) of every row of both banks are zero: a candidate whose rows start `off` rows
// below the first row of its group is the same 8 DP2A taps shifted by `off` bytes.
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

// |Hadamard| sum of the S x S tile whose column `d` this thread holds (S threads =
// S consecutive lanes); SAD when !SATD.  Must be called by all 32 lanes.
template <int S>
__device__ __forceinline__ uint32_t tile_dist(int (&d)[S], int lane, bool satd) {
  uint32_t acc = 0;
  if (satd) {
    distdev::hadamard_regs<S>(d);
    distdev::hadamard_lanes<S>(d, lane);
  }
#pragma unroll
  for (int i = 0; i < S; i++) acc += (uint32_t)abs(d[i]);
#pragma unroll
  for (int m = 1; m < S; m <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
  return acc;
}

// Two candidates of the same tile at once: x = d1 * 65536 + d0 is carried through the
// (linear) Hadamard as ONE integer per coefficient -- plain 32-bit adds / subs act on both
// halves, and since |coef| <= 64 * 255 < 2^15 the halves separate again exactly
// (lo = sign-extended low 16 bits, hi = (x - lo) >> 16).  Half the butterflies and shuffles
// per candidate.  Must be called by all 32 lanes.
template <int S>
__device__ __forceinline__ void tile_dist2(const int (&d0)[S], const int (&d1)[S], int lane, bool satd, uint32_t &a0,
                                           uint32_t &a1) {
  a0 = 0;
  a1 = 0;
  if (satd) {
    int x[S];
#pragma unroll
    for (int i = 0; i < S; i++) x[i] = d1[i] * 65536 + d0[i];
    distdev::hadamard_regs<S>(x);
    distdev::hadamard_lanes<S>(x, lane);
#pragma unroll
    for (int i = 0; i < S; i++) {
      const int lo = (int)(short)x[i];
      const int hi = (x[i] - lo) >> 16;
      a0 += (uint32_t)abs(lo);
      a1 += (uint32_t)abs(hi);
    }
  } else {
#pragma unroll
    for (int i = 0; i < S; i++) {
      a0 += (uint32_t)abs(d0[i]);
      a1 += (uint32_t)abs(d1[i]);
    }
  }
#pragma unroll
  for (int m = 1; m < S; m <<= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, m);
    a1 += __shfl_xor_sync(0xffffffffu, a1, m);
  }
}

struct Shared {
  const uint8_t *win, *org;
  uint32_t *dist;   // [NSLOT][8]
  Blk *blk;         // [NSLOT]
  const uint2 *taps;  // [2][16] packed s8 x 8
};

// slot -> (bsize, origin inside the superblock)
__device__ __forceinline__ void slot_info(int slot, int &bsize, int &rx, int &ry) {
  int off = 0;
  bsize = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    const int w = blk_w_of(b), h = blk_h_of(b);
    const int c = (w > 64 || h > 64) ? 0 : (64 / w) * (64 / h);
    if (slot < off + c) {
      bsize = b;
      break;
    }
    off += c;
  }
  const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
  const int k = slot - off;
  rx = (k % (64 / bw)) * bw;
  ry = (k / (64 / bw)) * bh;
}

// candidate k (0..7) of a telescopic step: (i, j) raster order without the centre
__device__ __forceinline__ r1c_mv step_cand(r1c_mv c, int step, int k) {
  const int kk = k < 4 ? k : k + 1;
  const int i = kk / 3, j = kk - 3 * i;
  r1c_mv mv;
  mv.row = (int16_t)(c.row + step * (i - 1));
  mv.col = (int16_t)(c.col + step * (j - 1));
  return mv;
}

// get_mv_range (src/me.rs:64-80) test for a block at absolute (bx, by)
__device__ __forceinline__ bool mv_in_range(const RefArgs &a, int bx, int by, int bw, int bh, r1c_mv mv) {
  const int xmin = -bx * 8 - 128 - 8 * bw, xmax = (a.W - bx - bw) * 8 + 128 + 8 * bw;
  const int ymin = -by * 8 - 128 - 8 * bh, ymax = (a.H - by - bh) * 8 + 128 + 8 * bh;
  return mv.col >= xmin && mv.col <= xmax && mv.row >= ymin && mv.row <= ymax;
}

// all tiles of one class for one step.  step == 0: the full-pel SATD re-evaluation
// (one candidate: the block's best_mv through the full-pel path).  A group of S threads
// takes a tile, decodes it and loads its org column ONCE, then walks the candidates
// (per-job set-up was a third of the instructions when every (tile, candidate) pair was
// a job of its own).
template <int S, int PB>
__device__ __forceinline__ void run_tiles(const RefArgs &a, const Shared &sh, const uint32_t *tiles, int t_lo,
                                          int t_hi, int step, int ncand, int RX, int RY, int wx0, int wy0) {
  constexpr int GROUPS = Rc<PB>::RT / S;
  // two candidates share one Hadamard as d1 * 65536 + d0 while |coef| < 2^15: always at 8 bits (64 * 255), at 10
  // bits only for the 4x4 tiles (16 * 1023)
  const bool PAIR = PB == 1 || (S == 4 && a.bit_depth <= 10);  // (a compile-time constant except for u16 4x4 tiles)
  // intermediate_bits = 4 up to 10 bits, 2 at 12 (src/mc.rs:222): with halved taps the two rounding shifts are
  // 3 - 1 and 11 - 1, or 5 - 1 and 9 - 1
  const int hsh = PB == 1 ? 2 : (a.bit_depth == 12 ? 4 : 2), hrnd = PB == 1 ? 2 : (1 << hsh) >> 1;
  const int vsh = PB == 1 ? 10 : (a.bit_depth == 12 ? 8 : 10), vrnd = PB == 1 ? 512 : (1 << vsh) >> 1;
  const int maxv = (1 << a.bit_depth) - 1;
  const int tid = threadIdx.x, lane = tid & 31, c = tid & (S - 1), grp = tid / S;
  const int nt = t_hi - t_lo;
  const int rounds = (nt + GROUPS - 1) / GROUPS;
  const bool satd = a.use_satd != 0;
  for (int rd = 0; rd < rounds; rd++) {
    const int t = rd * GROUPS + grp;
    bool tile_live = t < nt;
    int slot = 0, bw = 8, bh = 8, bx = 0, by = 0, tx = 0, ty = 0;
    int oc[S];
#pragma unroll
    for (int y = 0; y < S; y++) oc[y] = 0;
    r1c_mv center = {0, 0}, best = {0, 0};
    if (tile_live) {
      const uint32_t e = tiles[t_lo + t];
      slot = e & 1023;
      const int bsize = (e >> 10) & 31;
      bw = blk_w_of(bsize);
      bh = blk_h_of(bsize);
      bx = RX + 4 * ((e >> 15) & 15);
      by = RY + 4 * ((e >> 19) & 15);
      tx = 4 * ((e >> 23) & 15);
      ty = 4 * ((e >> 27) & 15);
      tile_live = a.off[bsize] >= 0 && bx + bw <= a.W && by + bh <= a.H;
      if (tile_live) {
        const Blk b = sh.blk[slot];
        center = b.center;
        best = b.mv;
        const int oi = (by - RY + ty) * 64 + (bx - RX + tx + c);
#pragma unroll
        for (int y = 0; y < S; y++)
          oc[y] = PB == 1 ? (int)sh.org[oi + y * 64] : (int)((const uint16_t *)sh.org)[oi + y * 64];
      }
    }
    const int tbx = (bw <= 4 ? 16 : 0), tby = (bh <= 4 ? 16 : 0);
    if (ncand == 1) {
      // step 0, full-pel path: ref block at po + mv / 8, truncating (src/me.rs:828-836)
      int d[S];
#pragma unroll
      for (int i = 0; i < S; i++) d[i] = 0;
      const bool live = tile_live && mv_in_range(a, bx, by, bw, bh, best);
      if (live) {
        int wx = bx + best.col / 8 + tx + c - wx0, wy = by + best.row / 8 + ty - wy0;
        wx = min(max(wx, 0), a.winW - 1);  // clamped for memory safety only
        wy = min(max(wy, 0), a.winH - S);
        const uint8_t *p = sh.win + wy * a.wstride + wx * PB;
#pragma unroll
        for (int y = 0; y < S; y++)
          d[y] = oc[y] - (PB == 1 ? (int)p[y * a.wstride] : (int)*(const uint16_t *)(p + y * a.wstride));
      }
      const uint32_t acc = tile_dist<S>(d, lane, satd);
      if (live && c == 0) atomicAdd(&sh.dist[slot * 8], acc);
    } else {
      // The 8 neighbours (i, j) of the centre, walked by column offset j: the (up to) three candidates of
      // one j have the same col_frac and source column, and their source rows start 0..2 rows apart, so
      // the S + 7 horizontally filtered rows are computed ONCE per j and every candidate is 4 DP2A per
      // output on the shared (row n, row n+1) pairs.  Candidates leave in pairs through the packed Hadamard.
      // get_mv_range test of the 3 candidate columns (bits 0..2) and rows (bits 3..5), once per tile
      uint32_t ok = 0;
      {
        const int xmin = -bx * 8 - 128 - 8 * bw, xmax = (a.W - bx - bw) * 8 + 128 + 8 * bw;
        const int ymin = -by * 8 - 128 - 8 * bh, ymax = (a.H - by - bh) * 8 + 128 + 8 * bh;
#pragma unroll
        for (int u = 0; u < 3; u++) {
          const int mc = (int)(int16_t)((int)center.col + step * (u - 1)), mr = (int)(int16_t)((int)center.row + step * (u - 1));
          ok |= (mc >= xmin && mc <= xmax ? 1u : 0u) << u;
          ok |= (mr >= ymin && mr <= ymax ? 8u : 0u) << u;
        }
        ok = tile_live ? ok : 0u;
      }
      int dp[S];  // the candidate waiting for its pair partner
      bool dp_live = false;
      int dp_k = 0, arrived = 0;  // (compile-time after unrolling)
      if (step == 8) {
        // First step: the centre is the full-pel winner, so all 8 neighbours are full-pel MVs and put_8tap copies
        // (src/mc.rs:224-231; the unified formula below gives the same pixels through unit taps: (16 px * 64 + 512) >> 10):
        // one window pixel per output, no filtering.  Same clamped origins as the general path.
#pragma unroll
        for (int jg = 0; jg < 3; jg++) {
          const int mvc = (int)center.col + 8 * (jg - 1);
          const int sx = max(min(bx + (mvc >> 3) - 3, a.ref.width), -a.ref.xorigin) + 3;
          const int wx = min(max(sx + tx + c - wx0, 0), a.winW - 1);  // clamped for memory safety only
#pragma unroll
          for (int i = 0; i < 3; i++) {
            if (i == 1 && jg == 1) continue;  // the centre itself
            const int kk = 3 * i + jg, k = kk < 4 ? kk : kk - 1;
            const int mvr = (int)center.row + 8 * (i - 1);
            const int sy = max(min(by + (mvr >> 3) - 3, a.ref.height), -a.ref.yorigin) + 3;
            const int wy = min(max(sy + ty - wy0, 0), a.winH - S);
            const uint8_t *p = sh.win + wy * a.wstride + wx * PB;
            const bool live = ((ok >> jg) & (ok >> (3 + i)) & 1u) != 0;
            int d[S];
#pragma unroll
            for (int t = 0; t < S; t++)
              d[t] = oc[t] - (PB == 1 ? (int)p[t * a.wstride] : (int)*(const uint16_t *)(p + t * a.wstride));
            if (!PAIR) {
              const uint32_t acc = tile_dist<S>(d, lane, satd);
              if (live && c == 0) atomicAdd(&sh.dist[slot * 8 + k], acc);
            } else if ((arrived++ & 1) == 0) {
#pragma unroll
              for (int t = 0; t < S; t++) dp[t] = d[t];
              dp_live = live;
              dp_k = k;
            } else {
              uint32_t a0, a1;
              tile_dist2<S>(dp, d, lane, satd, a0, a1);
              if (c == 0) {
                if (dp_live) atomicAdd(&sh.dist[slot * 8 + dp_k], a0);
                if (live) atomicAdd(&sh.dist[slot * 8 + k], a1);
              }
            }
          }
        }
        continue;
      }
      // NOT unrolled: the three column offsets unrolled are 3 x 9.5 KB of straight-line code in a kernel of 125 KB.
      // Alone on the GPU that was 3 % faster (1.24 against 1.28 ms per frame), but the lookahead window runs this kernel
      // next to me_sb2_kernel (103 KB) of the other stream on the same SMs, and there the smaller loop wins: 597 -> 614
      // Mpix/s for the window call (measured back to back, two runs each).  Rolling the step-8 loop above or the three
      // row offsets below as well lost on both counts.
#pragma unroll 1
      for (int jg = 0; jg < 3; jg++) {
        // predict_inter's get_params (src/predict.rs:267-284), luma, column part
        const int mvc = (int)center.col + step * (jg - 1);
        const int co = mvc >> 3, cf = (mvc - (co << 3)) << 1;
        const int sx = max(min(bx + co - 3, a.ref.width), -a.ref.xorigin) + 3;
        // window coordinates; clamped for memory safety only (a legal MV of the search range never
        // leaves the staged window)
        const int wx = min(max(sx + tx + c - wx0, 3), a.winW - 5);
        const int mvr0 = (int)center.row - step;
        const int sy0 = max(min(by + (mvr0 >> 3) - 3, a.ref.height), -a.ref.yorigin) + 3;
        const int wyb = min(max(sy0 + ty - wy0, 3), a.winH - S - 5);  // rows wyb - 2 .. wyb + S + 4
        uint32_t P[S + 6];  // P[n] = (mid[n + 1], mid[n + 2]), mid[n] = filtered window row wyb - 3 + n
        {
          const uint2 txw = sh.taps[tbx + cf];
          const uint8_t *row = sh.win + (wyb - 2) * a.wstride;
          int prev = hsum8<PB>(row, wx - 3, txw.x, txw.y, hrnd) >> hsh;
#pragma unroll
          for (int n = 0; n < S + 6; n++) {
            row += a.wstride;
            const int cur = hsum8<PB>(row, wx - 3, txw.x, txw.y, hrnd) >> hsh;
            P[n] = __byte_perm((uint32_t)prev, (uint32_t)cur, 0x5410);
            prev = cur;
          }
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
          if (i == 1 && jg == 1) continue;  // the centre itself
          const int kk = 3 * i + jg, k = kk < 4 ? kk : kk - 1;
          r1c_mv mv;
          mv.row = (int16_t)((int)center.row + step * (i - 1));
          mv.col = (int16_t)mvc;
          const int ro = (int)mv.row >> 3, rf = ((int)mv.row - (ro << 3)) << 1;
          const int sy = max(min(by + ro - 3, a.ref.height), -a.ref.yorigin) + 3;
          const uint32_t sh8 = 8u * (uint32_t)min(max(sy + ty - wy0 - wyb, 0), 2);
          const uint2 tyw = sh.taps[tby + rf];
          // taps 1..8 of the filter moved down by `off` rows (byte shifts of the 64-bit tap word)
          const uint32_t e0 = tyw.x << sh8, e1 = __funnelshift_l(tyw.x, tyw.y, sh8), e2 = __funnelshift_l(tyw.y, 0u, sh8);
          const uint32_t f0 = __funnelshift_r(e0, e1, 8), f1 = __funnelshift_r(e1, e2, 8);
          const bool live = ((ok >> jg) & (ok >> (3 + i)) & 1u) != 0;
          int d[S];  // of a candidate that is not live: computed all the same, never accumulated
#pragma unroll
          for (int t = 0; t < S; t++) {
            int v = dp2a_lo_ss(P[t], f0, vrnd);
            v = dp2a_hi_ss(P[t + 2], f0, v);
            v = dp2a_lo_ss(P[t + 4], f1, v);
            v = dp2a_hi_ss(P[t + 6], f1, v);
            d[t] = oc[t] - min(max(v >> vsh, 0), maxv);
          }
          if (!PAIR) {
            const uint32_t acc = tile_dist<S>(d, lane, satd);
            if (live && c == 0) atomicAdd(&sh.dist[slot * 8 + k], acc);
          } else if ((arrived++ & 1) == 0) {
#pragma unroll
            for (int t = 0; t < S; t++) dp[t] = d[t];
            dp_live = live;
            dp_k = k;
          } else {
            uint32_t a0, a1;
            tile_dist2<S>(dp, d, lane, satd, a0, a1);
            if (c == 0) {
              if (dp_live) atomicAdd(&sh.dist[slot * 8 + dp_k], a0);
              if (live) atomicAdd(&sh.dist[slot * 8 + k], a1);
            }
          }
        }
      }
    }
  }
}

__device__ unsigned char g_owner[NSLOT];  // which CTA of the superblock owns a slot

template <int PB>
__global__ void __launch_bounds__(Rc<PB>::RT, PB == 1 ? 2 : 1) me_refine_kernel(RefArgs a) {
  constexpr int RT = Rc<PB>::RT;
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x;
  const int sbi = blockIdx.x / SPLIT, part = blockIdx.x - sbi * SPLIT;
  const int RX = a.sbs[sbi].po_x, RY = a.sbs[sbi].po_y;
  uint8_t *s_win = smem;
  uint8_t *s_org = s_win + (((size_t)a.winH * a.wstride + 15) & ~(size_t)15);
  uint32_t *s_dist = (uint32_t *)(s_org + 64 * 64 * PB);
  Blk *s_blk = (Blk *)(s_dist + NSLOT * 8);
  uint2 *s_taps = (uint2 *)(s_blk + NSLOT);
  Shared sh = {s_win, s_org, s_dist, s_blk, s_taps};
  // window origin in plane coordinates
  const int wx0 = RX - a.range_x - MARGIN - 3, wy0 = RY - a.range_y - MARGIN - 3;

  // ---- stage: org superblock, reference window (clamped into the allocation), taps ----
  {
    const uint8_t *po = (const uint8_t *)a.org.p0, *pr = (const uint8_t *)a.ref.p0;
    const int oxmax = a.org.alloc_w - a.org.xorigin - 1, oymax = a.org.alloc_h - a.org.yorigin - 1;
    for (int i = tid; i < 64 * 64; i += RT) {
      const int y = min(max(RY + (i >> 6), -a.org.yorigin), oymax), x = min(max(RX + (i & 63), -a.org.xorigin), oxmax);
      if (PB == 1) s_org[i] = po[(long long)y * a.org.stride + x];
      else ((uint16_t *)s_org)[i] = ((const uint16_t *)po)[(long long)y * a.org.stride + x];
    }
    const int rxmax = a.ref.alloc_w - a.ref.xorigin - 1, rymax = a.ref.alloc_h - a.ref.yorigin - 1;
    const int wpr = a.wstride >> 2;
    for (int i = tid; i < a.winH * wpr; i += RT) {
      const int wy = i / wpr, w = i - wy * wpr;
      const int y = min(max(wy0 + wy, -a.ref.yorigin), rymax);
      const uint8_t *row = pr + (long long)y * a.ref.stride * PB;
      uint32_t v = 0;
#pragma unroll
      for (int b = 0; b < 4 / PB; b++) {
        const int x = min(max(wx0 + (4 / PB) * w + b, -a.ref.xorigin), rxmax);
        v |= (PB == 1 ? (uint32_t)row[x] : (uint32_t)((const uint16_t *)row)[x]) << (8 * PB * b);
      }
      ((uint32_t *)s_win)[i] = v;
    }
    if (tid < 32) {
      // halved taps; frac 0 = the unit tap 64 (see the unified formula above)
      const signed char *t = &c_taps[tid >> 4][tid & 15][0];
      uint32_t lo = 0, hi = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        lo |= (uint32_t)(uint8_t)(signed char)(t[k] / 2) << (8 * k);
        hi |= (uint32_t)(uint8_t)(signed char)(t[k + 4] / 2) << (8 * k);
      }
      if ((tid & 15) == 0) lo = 64u << 24, hi = 0;
      s_taps[tid] = make_uint2(lo, hi);
    }
  }
  // ---- per-block state from the full-pel results ------------------------------------------
  for (int slot = tid; slot < NSLOT; slot += RT) {
    Blk b;
    b.mv.row = b.mv.col = 0;
    b.center = b.mv;
    b.cost = ~0ull;
    if (g_owner[slot] == part) {
      int bsize, rx, ry;
      slot_info(slot, bsize, rx, ry);
      const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
      if (a.off[bsize] >= 0 && RX + rx + bw <= a.W && RY + ry + bh <= a.H) {
        const r1c_me_result r = a.results[a.off[bsize] + (long long)((RY + ry) / bh) * a.cols[bsize] + (RX + rx) / bw];
        b.mv = r.mv;
        b.center = r.mv;
        b.cost = r.cost;
      }
    }
    s_blk[slot] = b;
  }
  __syncthreads();

  const int nsteps = a.allow_hp ? 4 : 3;
  for (int s = a.use_satd ? -1 : 0; s < nsteps; s++) {
    const int step = s < 0 ? 0 : (8 >> s);
    const int ncand = s < 0 ? 1 : 8;
    for (int i = tid; i < NSLOT * 8; i += RT) s_dist[i] = 0;
    __syncthreads();
    run_tiles<8, PB>(a, sh, g_tiles8, a.tt.split8[part], a.tt.split8[part + 1], step, ncand, RX, RY, wx0, wy0);
    run_tiles<4, PB>(a, sh, g_tiles4, a.tt.split4[part], a.tt.split4[part + 1], step, ncand, RX, RY, wx0, wy0);
    __syncthreads();
    // ---- pick: strict `<` in candidate order (src/me.rs:934-937) ---------------------------
    for (int slot = tid; slot < NSLOT; slot += RT) {
      if (g_owner[slot] != part) continue;
      int bsize, rx, ry;
      slot_info(slot, bsize, rx, ry);
      const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
      const int bx = RX + rx, by = RY + ry;
      if (a.off[bsize] < 0 || bx + bw > a.W || by + bh > a.H) continue;
      Blk b = s_blk[slot];
      const int ln = (bw == 4 || bh == 4) ? 2 : 3;
      const r1c_mv zero = {0, 0};
      for (int k = 0; k < ncand; k++) {
        const r1c_mv mv = step == 0 ? b.mv : step_cand(b.center, step, k);
        unsigned long long cost = ~0ull;
        if (mv_in_range(a, bx, by, bw, bh, mv)) {
          uint32_t dist = s_dist[slot * 8 + k];
          if (a.use_satd) dist = (dist + ((1u << ln) >> 1)) >> ln;
          cost = 256ull * dist + (unsigned long long)me::mv_rate2(mv, zero, zero, a.allow_hp) * a.lambda;
        }
        if (step == 0) {
          b.cost = cost;  // lowest_cost = get_mv_rd_cost(best_mv), src/me.rs:235-252
        } else if (cost < b.cost) {
          b.cost = cost;
          b.mv = mv;
        }
      }
      b.center = b.mv;
      s_blk[slot] = b;
    }
    __syncthreads();
  }
  // ---- write back ------------------------------------------------------------------------------
  for (int slot = tid; slot < NSLOT; slot += RT) {
    if (g_owner[slot] != part) continue;
    int bsize, rx, ry;
    slot_info(slot, bsize, rx, ry);
    const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
    if (a.off[bsize] < 0 || RX + rx + bw > a.W || RY + ry + bh > a.H) continue;
    r1c_me_result r;
    r.mv = s_blk[slot].mv;
    r.reserved = 0;
    r.cost = s_blk[slot].cost;
    a.results[a.off[bsize] + (long long)((RY + ry) / bh) * a.cols[bsize] + (RX + rx) / bw] = r;
  }
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------
static void build_tile_tables(uint32_t *t4, uint32_t *t8, unsigned char *owner, TileTable *tt) {
  int n4 = 0, n8 = 0;
  // block boundaries (tile index where a block starts) to cut the lists at
  int starts4[MAX_T4 + 1], starts8[MAX_T8 + 1], ns4 = 0, ns8 = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    const int bw = blk_w_of(b), bh = blk_h_of(b);
    if (bw > 64 || bh > 64) continue;
    const bool small = bw == 4 || bh == 4;
    const int S = small ? 4 : 8;
    int k = 0;
    for (int y = 0; y < 64; y += bh)
      for (int x = 0; x < 64; x += bw, k++) {
        const int slot = slot_off(b) + k;
        if (small) starts4[ns4++] = n4;
        else starts8[ns8++] = n8;
        for (int ty = 0; ty < bh; ty += S)
          for (int tx = 0; tx < bw; tx += S) {
            const uint32_t e = (uint32_t)slot | ((uint32_t)b << 10) | ((uint32_t)(x / 4) << 15) | ((uint32_t)(y / 4) << 19) |
                               ((uint32_t)(tx / 4) << 23) | ((uint32_t)(ty / 4) << 27);
            if (small) t4[n4++] = e;
            else t8[n8++] = e;
          }
      }
  }
  starts4[ns4] = n4;
  starts8[ns8] = n8;
  tt->n4 = n4;
  tt->n8 = n8;
  auto cut = [](const int *starts, int ns, int target) {
    int best = starts[0];
    for (int i = 0; i <= ns; i++)
      if (abs(starts[i] - target) < abs(best - target)) best = starts[i];
    return best;
  };
  for (int c = 0; c <= SPLIT; c++) {
    tt->split4[c] = c == 0 ? 0 : (c == SPLIT ? n4 : cut(starts4, ns4, n4 * c / SPLIT));
    tt->split8[c] = c == 0 ? 0 : (c == SPLIT ? n8 : cut(starts8, ns8, n8 * c / SPLIT));
  }
  for (int c = 0; c < SPLIT; c++) {
    for (int i = tt->split4[c]; i < tt->split4[c + 1]; i++) owner[t4[i] & 1023] = (unsigned char)c;
    for (int i = tt->split8[c]; i < tt->split8[c + 1]; i++) owner[t8[i] & 1023] = (unsigned char)c;
  }
}

// Sub-pel tail of r1c_frame_me for 8-bit and 10-bit planes.  d_results: frame layout (off / cols of
// the mask), already holding the full-pel results of the superblocks in d_sbs.
// *handled = 0 -> not applicable, the caller uses the generic path (subpel.cu).
int r1c_refine_frame(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_frame_me_params *p,
                     int bit_depth, const long long *off, const int *cols, const int *rows, const r1c_sb_job *d_sbs, int64_t n_sb,
                     r1c_me_result *d_results, int *handled) {
  *handled = 0;
  const char *force = getenv("R1C_REFINE_GENERIC");
  const int pb = org->px_bytes;
  if ((force && force[0] == '1') || pb != ref->px_bytes) return R1C_OK;
  RefArgs a;
  a.winW = 64 + 2 * (p->range_x + MARGIN) + 7;
  a.winH = 64 + 2 * (p->range_y + MARGIN) + 7;
  a.wstride = ((a.winW * pb + 3) & ~3) + 8;  // bytes
  a.bit_depth = bit_depth;
  const size_t smem = (((size_t)a.winH * a.wstride + 15) & ~(size_t)15) + 64 * 64 * pb + (size_t)NSLOT * 8 * 4 + (size_t)NSLOT * sizeof(Blk) + 32 * 8;
  if (smem > (pb == 1 ? 110 : 220) * 1024) return R1C_OK;
  *handled = 1;
  if (n_sb == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  static TileTable tt;
  static bool built = false, uploaded[64] = {false};
  static uint32_t t4[MAX_T4], t8[MAX_T8];
  static unsigned char owner[NSLOT];
  if (!built) {
    build_tile_tables(t4, t8, owner, &tt);
    built = true;
  }
  if (!uploaded[ctx->device & 63]) {
    R1C_CUDA(ctx, cudaMemcpyToSymbol(g_tiles4, t4, sizeof t4));
    R1C_CUDA(ctx, cudaMemcpyToSymbol(g_tiles8, t8, sizeof t8));
    R1C_CUDA(ctx, cudaMemcpyToSymbol(g_owner, owner, sizeof owner));
    uploaded[ctx->device & 63] = true;
  }
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.sbs = d_sbs;
  a.results = d_results;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    a.off[b] = off[b];
    a.cols[b] = cols[b];
    a.rows[b] = rows[b];
  }
  a.tt = tt;
  a.W = p->width;
  a.H = p->height;
  a.range_x = p->range_x;
  a.range_y = p->range_y;
  a.lambda = p->lambda;
  a.use_satd = p->use_satd;
  a.allow_hp = p->allow_high_precision_mv;
  auto kern = pb == 1 ? me_refine_kernel<1> : me_refine_kernel<2>;
  R1C_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  r1c_prof_begin(ctx, R1C_PROF_SUBPEL_COST);
  kern<<<(unsigned)(n_sb * SPLIT), pb == 1 ? Rc<1>::RT : Rc<2>::RT, smem, ctx->stream>>>(a);
  r1c_prof_end(ctx, R1C_PROF_SUBPEL_COST);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}
