// me_sb2.cu -- whole-superblock full search, second generation (8-bit planes and u16 planes: up to 10 bits, and 12 bits).
//
// Same contract as me_sb.cu (the result of the reference's full_search,
// src/me.rs:943-990, for every block of every size inside a 64x64 region against
// FullSearch::full_pixel_me's window, src/me.rs:533-546), restructured around
// what bounds it on sm_100a: the integer ALU pipe and shared-memory loads.
//
//  * Work item = (superblock, chunk of candidate rows); 512 threads: warp w owns
//    the 16x16 quadrant w, lane l owns one SLOT per iteration.  A slot is a pair
//    of vertically adjacent candidates (dx, dy) / (dx, dy + 1): they share 15 of
//    their 16 reference rows and every org row, so the pair costs 17 x 4 LDS.32
//    + 16 LDS.128 for 128 VABSDIFF4.
//  * The reference window is staged four times, shifted by 0..3 bytes, so every
//    16-byte row segment of every candidate starts on a word: no funnel shifts
//    in the inner loop (SHF shares the half-rate ALU pipe with VABSDIFF4).
//  * Running minimum of the 44 blocks up to 8x8 / 4x16 / 16x4 (SAD < 2^14) is ONE
//    32-bit key  (sad + rl >> 8) << 17 | (rl & 255) << 9 | seq  -> IMAD + VIMNMX
//    per update; `seq` numbers the thread's candidates in raster order, so the
//    minimum key is the FIRST minimum of the reference's strict-`<` raster scan.
//    The larger blocks keep (cost32, seq) pairs.  rl = rate * lambda.
//  * Blocks spanning quadrants: five partial sums per (quadrant, candidate) go
//    through a double-buffered shared-memory exchange (one barrier / iteration).
//  * Each work item writes 849 partial keys  cost << 32 | raster index;
//    sb2_finalize_kernel takes the minimum over the chunks (lexicographic =
//    raster tie-break), turns it into (mv, cost) and scatters it either to the
//    [job][849] layout of r1c_me_sb_batch or straight to a caller-supplied map
//    (whole-frame layout of r1c_frame_me).
//  * u16 planes (template PB = 2, bit depth <= 10): two pixels per window word, two pixel
//    phases; sum |a - b| = 2 sum max(a, b) - sum a - sum b: the inner loop accumulates packed
//    maxima only (VIMNMX.U16x2, 8 values per 16-bit half), sum b of a cell is a box filter of
//    the window (negated s16 table in shared memory), sum a of a block leaves the cost in
//    sb2_finalize_kernel (see max2 / fold_cell).  An 8x8 SAD + sum a needs 17 bits, so the key
//    keeps 6 bits of seq (12 candidate rows per chunk, 11 chunks for +-64) and 18 bits of
//    sad + rl >> 8; (SAD + sum a) * 256 + rl of a 64x64 block still fits 32 bits (not at 12 bits).
//  * MV rates come from per-column / per-row tables in shared memory; in border superblocks
//    only the candidates that are illegal for some block pay for the get_mv_range masks.
// INT-bound: per (pixel, candidate) 1/4 VABSDIFF4 (u8) or 3/4 full-rate instructions (u16);
// planes are read once per work item from L2.  12-bit planes: the D12 variant (Px).  Falls back to me_sb.cu for
// step 2, huge lambda or windows too small / too large for this layout.
#include <stdlib.h>

#include "me_common.cuh"

namespace {

constexpr int T2 = 512;
// per pixel size: u8 planes (PB 1) and 10-bit u16 planes (PB 2)
template <int PB, bool D12 = false>
struct Px {
  // The 8x8 SAD of the packed-key blocks is < 2^14 at 8 bits but needs 16 bits at 10: the 10-bit kernel has 6
  // bits of `seq` (candidate rows are cut into more chunks) and 18 bits of sad + rl >> 8.
  // 12 bits (D12): the cells hold the TRUE SAD (sum org of the cell comes off in the search loop, 16 registers), an
  // 8x8 SAD is < 2^18, the field has 19 bits and `seq` 5 (6 candidate rows per chunk, 22 chunks for +-64); 256 * SAD
  // + rl of a 64x64 block stays below 2^32 - 1 for lambda <= 16000 (sb2_plan).
  static constexpr int SEQ_BITS = PB == 1 ? 9 : (D12 ? 5 : 6);
  static constexpr int KSHIFT = 8 + SEQ_BITS;  // sad field of the 32-bit key
  static constexpr int PXW = 4 / PB;           // pixels per window word = phases of the staged window
  static constexpr int NWQ = 4 * PB;           // words of a 16-pixel quadrant row
};
constexpr uint32_t NONE32 = 0xffffffffu;
constexpr int NSLOT = 849;

struct Sb2Args {
  PlaneView org, ref;
  r1c_sb_params p;
  const r1c_sb_job *jobs;
  unsigned long long *partial;  // [n * split][849]
  uint32_t *bias;               // u16 planes: [n][849] sum of org over each block (written by chunk 0)
  long long n;
  int split;  // chunks of candidate rows per superblock
  int rpc;    // candidate rows per chunk (even)
  int rsw;    // words per staged window row
  int csz;    // words of the staged window: rows * rsw * 4 byte phases
  int M;      // ceil(NX / 32)
};

__host__ __device__ inline int sb2_count(int b) {
  const int w = blk_w_of(b), h = blk_h_of(b);
  return (w > 64 || h > 64) ? 0 : (64 / w) * (64 / h);
}
// first slot of BlockSize b in the [849] layout = sum of sb2_count over the smaller enum values
__constant__ short c_sb2_off[R1C_BLOCK_SIZES_ALL] = {0,   256, 384, 512, 576, 608, 640, 656, 664, 672, 676,
                                                     678, 680, 681, 681, 681, 681, 745, 809, 825, 841, 845};
__host__ __device__ inline int sb2_offset(int b) {
#ifdef __CUDA_ARCH__
  return c_sb2_off[b];
#else
  int o = 0;
  for (int i = 0; i < b; i++) o += sb2_count(i);
  return o;
#endif
}

__device__ __forceinline__ uint32_t rate1(int diff, int allow_hp) {
  const int d = allow_hp ? (int)(int16_t)diff : ((int)(int16_t)diff >> 1);
  if (d == 0) return 0;
  const uint32_t a = (uint32_t)(d < 0 ? -d : d) & 0xffffu;
  return 2u * (uint32_t)(32 - __clz(a));
}

struct Geo2 {
  int QX, QY;  // absolute position of the warp's quadrant
  int RX, RY;  // region origin
  int W4, H4;
};
// legality of MV offset (ox, oy) for a block at (bx, by) of size (bw, bh): the
// window of full_pixel_me after get_mv_range (src/me.rs:64-80, 533-546)
__device__ __forceinline__ bool mv_ok(const Geo2 &g, int bx, int by, int bw, int bh, int ox, int oy) {
  return (bx + ox + bw >= -16) && (bx + ox <= g.W4 + 16) && (by + oy + bh >= -16) && (by + oy <= g.H4 + 16);
}

// every block of the 64x64 region may take MV offset (ox, oy): the 4x4 blocks on the region's edges decide
__device__ __forceinline__ bool cand_free(const Geo2 &g, int ox, int oy) {
  return (g.RX + 4 + ox >= -16) && (g.RX + 60 + ox <= g.W4 + 16) && (g.RY + 4 + oy >= -16) && (g.RY + 60 + oy <= g.H4 + 16);
}

// in-quadrant aggregation + running minima of one candidate
template <bool MASKED, int KSHIFT>
__device__ __forceinline__ void quad_update(const uint32_t (&cell)[4][4], uint32_t rck, uint32_t rl, uint32_t seq,
                                            uint32_t (&best)[44], uint32_t (&pc)[5], uint32_t (&ps)[5],
                                            uint32_t (&xs)[5], const Geo2 &g, int ox, int oy) {
#define K_UPD(BI, SAD, LX, LY, BW, BH)                                                     \
  do {                                                                                     \
    uint32_t key_ = (SAD) * (1u << KSHIFT) + rck;                                          \
    if (MASKED) key_ = mv_ok(g, g.QX + (LX), g.QY + (LY), BW, BH, ox, oy) ? key_ : NONE32; \
    best[BI] = min(best[BI], key_);                                                        \
  } while (0)
#define P_UPD(PI, SAD, LX, LY, BW, BH)                                                  \
  do {                                                                                  \
    uint32_t c_ = (SAD) * 256u + rl;                                                    \
    if (MASKED) c_ = mv_ok(g, g.QX + (LX), g.QY + (LY), BW, BH, ox, oy) ? c_ : NONE32;  \
    if (c_ < pc[PI]) {                                                                  \
      pc[PI] = c_;                                                                      \
      ps[PI] = seq;                                                                     \
    }                                                                                   \
  } while (0)
  uint32_t v48[2][4], h84[4][2], s88[2][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) K_UPD(i * 4 + k, cell[i][k], 4 * k, 4 * i, 4, 4);
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) {
      v48[i][k] = cell[2 * i][k] + cell[2 * i + 1][k];
      K_UPD(16 + i * 4 + k, v48[i][k], 4 * k, 8 * i, 4, 8);
    }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int k = 0; k < 2; k++) {
      h84[i][k] = cell[i][2 * k] + cell[i][2 * k + 1];
      K_UPD(24 + i * 2 + k, h84[i][k], 8 * k, 4 * i, 8, 4);
    }
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int k = 0; k < 2; k++) {
      s88[i][k] = v48[i][2 * k] + v48[i][2 * k + 1];
      K_UPD(32 + i * 2 + k, s88[i][k], 8 * k, 8 * i, 8, 8);
    }
#pragma unroll
  for (int k = 0; k < 4; k++) K_UPD(36 + k, v48[0][k] + v48[1][k], 4 * k, 0, 4, 16);
#pragma unroll
  for (int i = 0; i < 4; i++) K_UPD(40 + i, h84[i][0] + h84[i][1], 0, 4 * i, 16, 4);
  const uint32_t c816l = s88[0][0] + s88[1][0], c816r = s88[0][1] + s88[1][1];
  const uint32_t c168t = s88[0][0] + s88[0][1], c168b = s88[1][0] + s88[1][1];
  const uint32_t c1616 = c816l + c816r;
  P_UPD(0, c816l, 0, 0, 8, 16);
  P_UPD(1, c816r, 8, 0, 8, 16);
  P_UPD(2, c168t, 0, 0, 16, 8);
  P_UPD(3, c168b, 0, 8, 16, 8);
  P_UPD(4, c1616, 0, 0, 16, 16);
  xs[0] = c1616;
  xs[1] = c816l;
  xs[2] = c816r;
  xs[3] = c168t;
  xs[4] = c168b;
#undef K_UPD
#undef P_UPD
}

template <bool V>
struct BoolTag {
  static constexpr bool value = V;
};

struct Chunk {
  int NX, NYc, NP, NS, iters;
  int row0;        // first candidate row of the chunk
  int cxlo, cylo;  // MV offset of candidate (0, 0) of the whole window
};

// running minimum of one block spanning quadrants, for this lane's candidates A and B
template <bool MASKED>
__device__ __forceinline__ void big_upd(uint32_t &bc, uint32_t &bs, uint32_t sA, uint32_t sB, uint32_t rlA,
                                        uint32_t rlB, uint32_t seqA, uint32_t seqB, const Geo2 &g, int lx, int ly,
                                        int bw, int bh, int ox, int oyA) {
  uint32_t cA = sA * 256u + rlA, cB = sB * 256u + rlB;
  if (MASKED) {
    cA = mv_ok(g, g.RX + lx, g.RY + ly, bw, bh, ox, oyA) ? cA : NONE32;
    cB = mv_ok(g, g.RX + lx, g.RY + ly, bw, bh, ox, oyA + 1) ? cB : NONE32;
  }
  const bool la = cA < bc;
  bc = la ? cA : bc;
  bs = la ? seqA : bs;
  const bool lb = cB < bc;
  bc = lb ? cB : bc;
  bs = lb ? seqB : bs;
}

// The 65 blocks that span quadrants, split over the 16 warps by role (q = warp):
//   q 0..3   column c = q      : 16x32 top / bottom, 16x64, and the 32x32 block c
//   q 4..7   column c = q - 4  : 8x32 left-top, left-bottom, right-top, right-bottom
//   q 8..11  row r = q - 8     : 32x16 left / right, 64x16, and 32x64 left / right (r = 0, 1)
//                                or 64x32 top / bottom (r = 2, 3)
//   q 12..15 row r = q - 12    : 32x8 top-left, top-right, bottom-left, bottom-right;
//                                q 12 also the 64x64
// slot s of role q <-> (bsize, lx, ly): big_slot().  X(part, quadrant) = exchanged sums.
struct BigSlot {
  int bsize, lx, ly;
};
__host__ __device__ inline BigSlot big_slot(int q, int s) {
  BigSlot b = {-1, 0, 0};
  if (q < 4) {
    const int c = q;
    if (s == 0) b = {R1C_BLOCK_16X32, 16 * c, 0};
    else if (s == 1) b = {R1C_BLOCK_16X32, 16 * c, 32};
    else if (s == 2) b = {R1C_BLOCK_16X64, 16 * c, 0};
    else if (s == 3) b = {R1C_BLOCK_32X32, 32 * (c & 1), 32 * (c >> 1)};
  } else if (q < 8) {
    const int c = q - 4;
    if (s < 4) b = {R1C_BLOCK_8X32, 16 * c + 8 * (s >> 1), 32 * (s & 1)};
  } else if (q < 12) {
    const int r = q - 8;
    if (s == 0) b = {R1C_BLOCK_32X16, 0, 16 * r};
    else if (s == 1) b = {R1C_BLOCK_32X16, 32, 16 * r};
    else if (s == 2) b = {R1C_BLOCK_64X16, 0, 16 * r};
    else if (s == 3) {
      if (r < 2) b = {R1C_BLOCK_32X64, 32 * r, 0};
      else b = {R1C_BLOCK_64X32, 0, 32 * (r - 2)};
    }
  } else {
    const int r = q - 12;
    if (s < 4) b = {R1C_BLOCK_32X8, 32 * (s & 1), 16 * r + 8 * (s >> 1)};
    else if (s == 4 && q == 12) b = {R1C_BLOCK_64X64, 0, 0};
  }
  return b;
}

// u16 planes: sum |a - b| = 2 sum max(a, b) - sum a - sum b.  The inner loop only accumulates the packed maxima
// (VIMNMX.U16x2 + a 3-input add per two words); sum b of a 4x4 cell at a window position is candidate-independent
// work (a box filter of the staged window, kept negated as s16 in shared memory), and sum a of a block is the same
// for all its candidates, so the running minima are taken on SAD + sum a and the constant leaves in the epilogue.
__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
// 2 * (lo16(c) + hi16(c)) + nb
__device__ __forceinline__ uint32_t fold_cell(uint32_t c, int nb) {
  uint32_t d;
  asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(c), "r"(0x0202u), "r"((uint32_t)nb));
  return d;
}

template <bool MASKED, int RSW, int PB, bool D12>
__device__ __forceinline__ void sb2_search(const Sb2Args &a, const r1c_sb_job &job, const Chunk &ck,
                                           const uint32_t *s_org, const uint32_t *s_ref, uint32_t *s_x,
                                           const int16_t *s_nbox, const uint32_t *s_cell, const uint32_t *s_rcol, const uint32_t *s_rrow,
                                           uint32_t (&best)[44], uint32_t (&pc)[5], uint32_t (&ps)[5],
                                           uint32_t (&bc)[5], uint32_t (&bs)[5], const Geo2 &g) {
  const int tid = threadIdx.x, lane = tid & 31, q = tid >> 5;
  const int qx = q & 3, qy = q >> 2;
  constexpr int SEQ_BITS = Px<PB, D12>::SEQ_BITS, KSHIFT = Px<PB, D12>::KSHIFT, PXW = Px<PB>::PXW, NWQ = Px<PB>::NWQ;
  const int rsw = RSW ? RSW : a.rsw;
  const int sdx = 32 % ck.NX, sj = 32 / ck.NX;
  // 12 bits: sum org of this quadrant's 16 cells + the offset of the box table (see the kernel), taken off every cell
  uint32_t sa[4][4];
  if (D12) {
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) sa[i][k] = s_cell[(qy * 4 + i) * 16 + qx * 4 + k] + 32760u;
  }
  const int M2 = 2 * a.M;
  // slot of this lane: f = it * 32 + lane = j * NX + dx
  int f = lane, j = lane / ck.NX, dx = lane - j * ck.NX;
  const uint32_t *orow = s_org + (qy * 16) * (16 * PB) + qx * NWQ;
  // role of this warp in the cross-quadrant phase
  const int rg = q >> 2, ri = q & 3;
  for (int it = 0; it < ck.iters; it++) {
    const int dyA = min(2 * j, ck.NYc - 2);
    // window word (row, w) of pixel phase p lives at s_ref[(row * rsw + w) * PXW + p]
    const uint32_t *rp = s_ref + ((qy * 16 + dyA) * rsw + qx * NWQ + dx / PXW) * PXW + dx % PXW;
    // ---- 2 x 16 cell SADs: candidate A = (dx, dyA), B = (dx, dyA + 1) ----------------
    uint32_t cA[4][4], cB[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) cA[i][k] = cB[i][k] = 0;
    if (PB == 1) {
      uint4 oprev = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int r = 0; r < 17; r++) {
        const uint32_t *rr = rp + r * rsw * 4;
        const uint32_t w0 = rr[0], w1 = rr[4], w2 = rr[8], w3 = rr[12];
        if (r >= 1) {
          const int i = (r - 1) >> 2;
          cB[i][0] = me::sad4_acc(oprev.x, w0, cB[i][0]);
          cB[i][1] = me::sad4_acc(oprev.y, w1, cB[i][1]);
          cB[i][2] = me::sad4_acc(oprev.z, w2, cB[i][2]);
          cB[i][3] = me::sad4_acc(oprev.w, w3, cB[i][3]);
        }
        if (r < 16) {
          const uint4 o = *(const uint4 *)(orow + r * 16);
          const int i = r >> 2;
          cA[i][0] = me::sad4_acc(o.x, w0, cA[i][0]);
          cA[i][1] = me::sad4_acc(o.y, w1, cA[i][1]);
          cA[i][2] = me::sad4_acc(o.z, w2, cA[i][2]);
          cA[i][3] = me::sad4_acc(o.w, w3, cA[i][3]);
          oprev = o;
        }
      }
    } else {
      // u16 pixels: two per word, packed maxima (VIMNMX.U16x2 + add, see max2); a cell accumulates 8 values
      // in each 16-bit half (<= 8 * 1023) and is folded after the loop.  The quadrant row is taken
      // in two 8-pixel halves to keep the live registers down; org words are the same address for the
      // whole warp (broadcast), so B re-reads its row instead of keeping it.
      uint4 oprev[2];
#pragma unroll
      for (int r = 0; r < 17; r++) {
#pragma unroll
        for (int hf = 0; hf < 2; hf++) {
          const uint32_t *rr = rp + (r * rsw + 4 * hf) * 2;
          const uint32_t w0 = rr[0], w1 = rr[2], w2 = rr[4], w3 = rr[6];
          if (r >= 1) {
            const uint4 o = oprev[hf];
            const int i = (r - 1) >> 2;
            cB[i][2 * hf] += max2(o.x, w0) + max2(o.y, w1);
            cB[i][2 * hf + 1] += max2(o.z, w2) + max2(o.w, w3);
          }
          if (r < 16) {
            const uint4 o = *(const uint4 *)(orow + r * 32 + 4 * hf);
            const int i = r >> 2;
            cA[i][2 * hf] += max2(o.x, w0) + max2(o.y, w1);
            cA[i][2 * hf + 1] += max2(o.z, w2) + max2(o.w, w3);
            oprev[hf] = o;
          }
        }
      }
      // cell = 2 * sum max - box sum of the window at (candidate, cell) = SAD + sum org of the cell
      const int bw = rsw * PXW;
      const int16_t *nb = s_nbox + (qy * 16 + dyA) * bw + qx * 16 + dx;
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int k = 0; k < 4; k++) {
          cA[i][k] = fold_cell(cA[i][k], nb[4 * i * bw + 4 * k]);
          cB[i][k] = fold_cell(cB[i][k], nb[(4 * i + 1) * bw + 4 * k]);
          if (D12) {  // the true SAD of the cell
            cA[i][k] -= sa[i][k];
            cB[i][k] -= sa[i][k];
          }
        }
    }
    // ---- rate terms ---------------------------------------------------------------------
    const int ox = ck.cxlo + dx, oyA = ck.cylo + ck.row0 + dyA;
    // rate = min(rate(mv, pmv0), rate(mv, pmv1) + 1) from the per-column / per-row tables (two u16 halves each)
    const uint32_t rcp = s_rcol[dx], rsA = rcp + s_rrow[dyA], rsB = rcp + s_rrow[dyA + 1];
    const uint32_t rA = min(rsA & 0xffffu, rsA >> 16), rB = min(rsB & 0xffffu, rsB >> 16);
    const uint32_t rlA = rA * a.p.lambda, rlB = rB * a.p.lambda;
    const uint32_t seqA = (uint32_t)(j * M2 + (dx >> 5)), seqB = seqA + (uint32_t)a.M;
    const uint32_t rckA = ((rlA >> 8) << KSHIFT) | ((rlA & 255u) << SEQ_BITS) | seqA;
    const uint32_t rckB = ((rlB >> 8) << KSHIFT) | ((rlB & 255u) << SEQ_BITS) | seqB;

    uint32_t xsA[5], xsB[5];
    // border superblocks: most candidates are still legal for every block of the region and take the unmasked code
    const bool freeA = !MASKED || cand_free(g, ox, oyA), freeB = !MASKED || cand_free(g, ox, oyA + 1);
    if (freeA) quad_update<false, KSHIFT>(cA, rckA, rlA, seqA, best, pc, ps, xsA, g, ox, oyA);
    else quad_update<MASKED, KSHIFT>(cA, rckA, rlA, seqA, best, pc, ps, xsA, g, ox, oyA);
    if (freeB) quad_update<false, KSHIFT>(cB, rckB, rlB, seqB, best, pc, ps, xsB, g, ox, oyA + 1);
    else quad_update<MASKED, KSHIFT>(cB, rckB, rlB, seqB, best, pc, ps, xsB, g, ox, oyA + 1);

    // ---- exchange of the quadrant partial sums: xb[part][quadrant][A lanes | B lanes] ---
    uint32_t *xb = s_x + (it & 1) * (5 * 16 * 64);
#pragma unroll
    for (int p = 0; p < 5; p++) {
      xb[(p * 16 + q) * 64 + lane] = xsA[p];
      xb[(p * 16 + q) * 64 + 32 + lane] = xsB[p];
    }
    __syncthreads();
    // ---- blocks spanning quadrants, by warp role ----------------------------------------
    auto big_phase = [&](auto masked_tag) {
      constexpr bool BM = decltype(masked_tag)::value;
      const uint32_t *xl = xb + lane;
#define XA(PART, QUAD) xl[((PART) * 16 + (QUAD)) * 64]
#define XB(PART, QUAD) xl[((PART) * 16 + (QUAD)) * 64 + 32]
#define BIG(S, SA, SB, LX, LY, BW, BH) \
  big_upd<BM>(bc[S], bs[S], SA, SB, rlA, rlB, seqA, seqB, g, LX, LY, BW, BH, ox, oyA)
      if (rg == 0) {  // column ri: 16x16 sums of its four quadrants
        const int c = ri;
        const uint32_t a0 = XA(0, c), a1 = XA(0, c + 4), a2 = XA(0, c + 8), a3 = XA(0, c + 12);
        const uint32_t b0 = XB(0, c), b1 = XB(0, c + 4), b2 = XB(0, c + 8), b3 = XB(0, c + 12);
        const uint32_t at = a0 + a1, ab = a2 + a3, bt = b0 + b1, bb = b2 + b3;
        BIG(0, at, bt, 16 * c, 0, 16, 32);
        BIG(1, ab, bb, 16 * c, 32, 16, 32);
        BIG(2, at + ab, bt + bb, 16 * c, 0, 16, 64);
        const int q0 = (c >> 1) * 8 + (c & 1) * 2;  // 32x32 block c
        const uint32_t sA = XA(0, q0) + XA(0, q0 + 1) + XA(0, q0 + 4) + XA(0, q0 + 5);
        const uint32_t sB = XB(0, q0) + XB(0, q0 + 1) + XB(0, q0 + 4) + XB(0, q0 + 5);
        BIG(3, sA, sB, 32 * (c & 1), 32 * (c >> 1), 32, 32);
      } else if (rg == 1) {  // column ri: 8x16 halves -> 8x32
        const int c = ri;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const uint32_t a0 = XA(1 + h, c), a1 = XA(1 + h, c + 4), a2 = XA(1 + h, c + 8), a3 = XA(1 + h, c + 12);
          const uint32_t b0 = XB(1 + h, c), b1 = XB(1 + h, c + 4), b2 = XB(1 + h, c + 8), b3 = XB(1 + h, c + 12);
          BIG(2 * h, a0 + a1, b0 + b1, 16 * c + 8 * h, 0, 8, 32);
          BIG(2 * h + 1, a2 + a3, b2 + b3, 16 * c + 8 * h, 32, 8, 32);
        }
      } else if (rg == 2) {  // row ri
        const int r = ri;
        const uint32_t a0 = XA(0, 4 * r), a1 = XA(0, 4 * r + 1), a2 = XA(0, 4 * r + 2), a3 = XA(0, 4 * r + 3);
        const uint32_t b0 = XB(0, 4 * r), b1 = XB(0, 4 * r + 1), b2 = XB(0, 4 * r + 2), b3 = XB(0, 4 * r + 3);
        const uint32_t al = a0 + a1, ar = a2 + a3, bl = b0 + b1, br = b2 + b3;
        BIG(0, al, bl, 0, 16 * r, 32, 16);
        BIG(1, ar, br, 32, 16 * r, 32, 16);
        BIG(2, al + ar, bl + br, 0, 16 * r, 64, 16);
        // r < 2: 32x64 half r (quadrant columns 2r, 2r+1); else 64x32 half r-2 (rows 2(r-2), +1)
        const int q0 = r < 2 ? 2 * r : 8 * (r - 2), st = r < 2 ? 4 : 1, st2 = r < 2 ? 1 : 4;
        uint32_t sA = 0, sB = 0;
#pragma unroll
        for (int t = 0; t < 4; t++) {
          sA += XA(0, q0 + t * st) + XA(0, q0 + t * st + st2);
          sB += XB(0, q0 + t * st) + XB(0, q0 + t * st + st2);
        }
        BIG(3, sA, sB, r < 2 ? 32 * r : 0, r < 2 ? 0 : 32 * (r - 2), r < 2 ? 32 : 64, r < 2 ? 64 : 32);
      } else {  // row ri: 16x8 halves -> 32x8
        const int r = ri;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const uint32_t a0 = XA(3 + h, 4 * r), a1 = XA(3 + h, 4 * r + 1), a2 = XA(3 + h, 4 * r + 2), a3 = XA(3 + h, 4 * r + 3);
          const uint32_t b0 = XB(3 + h, 4 * r), b1 = XB(3 + h, 4 * r + 1), b2 = XB(3 + h, 4 * r + 2), b3 = XB(3 + h, 4 * r + 3);
          BIG(2 * h, a0 + a1, b0 + b1, 0, 16 * r + 8 * h, 32, 8);
          BIG(2 * h + 1, a2 + a3, b2 + b3, 32, 16 * r + 8 * h, 32, 8);
        }
        if (r == 0) {  // 64x64
          uint32_t sA = 0, sB = 0;
#pragma unroll
          for (int t = 0; t < 16; t++) {
            sA += XA(0, t);
            sB += XB(0, t);
          }
          BIG(4, sA, sB, 0, 0, 64, 64);
        }
      }
#undef XA
#undef XB
#undef BIG
    };
    if (freeA && freeB) big_phase(BoolTag<false>{});
    else big_phase(BoolTag<MASKED>{});
    // ---- next slot; a lane that runs out of slots repeats its last one (same keys) ----
    if (f + 32 < ck.NS) {
      f += 32;
      dx += sdx;
      j += sj;
      if (dx >= ck.NX) {
        dx -= ck.NX;
        j++;
      }
    }
  }
}

// (seq, lane) -> raster index of the candidate inside the WHOLE window
__device__ __forceinline__ uint32_t decode_ridx(const Chunk &ck, uint32_t seq, int lane, int M, uint32_t inv_m2) {
  const int j = (int)((seq * inv_m2) >> 16);  // seq / (2M): exact for seq < 512, 2M <= 16
  const int rem = (int)seq - j * 2 * M;
  const int isB = rem >= M, k = rem - (isB ? M : 0);
  const int d0 = (lane - j * ck.NX) & 31;
  const int dx = d0 + 32 * k;
  const int dy = min(2 * j, ck.NYc - 2) + isB;
  return (uint32_t)((ck.row0 + dy) * ck.NX + dx);
}

// lexicographic (cost, raster index) minimum over the warp -> 64-bit partial key
__device__ __forceinline__ unsigned long long warp_best(uint32_t cost, uint32_t ridx) {
  const uint32_t cmin = __reduce_min_sync(0xffffffffu, cost);
  const uint32_t rmin = __reduce_min_sync(0xffffffffu, cost == cmin ? ridx : NONE32);
  return cmin == NONE32 ? ~0ull : (((unsigned long long)cmin << 32) | rmin);
}

template <int RSW, int PB, bool D12 = false>
__global__ void __launch_bounds__(T2, 1) me_sb2_kernel(Sb2Args a) {
  static_assert(!D12 || PB == 2, "12-bit planes are u16");
  constexpr int SEQ_BITS = Px<PB, D12>::SEQ_BITS, PXW = Px<PB>::PXW;
  extern __shared__ __align__(16) unsigned char smem[];
  const long long wi = blockIdx.x;
  const long long ji = wi / a.split;
  const int chunk = (int)(wi - ji * a.split);
  const r1c_sb_job job = a.jobs[ji];
  const int tid = threadIdx.x, lane = tid & 31, q = tid >> 5;
  const int Rx = a.p.range_x, Ry = a.p.range_y;
  const int cx = job.cmv.col / 8, cy = job.cmv.row / 8;  // truncating, like the reference
  const int rsw = RSW ? RSW : a.rsw;

  Chunk ck;
  ck.NX = 2 * Rx + 1;
  const int NY = 2 * Ry + 1;
  ck.row0 = chunk * a.rpc;
  ck.NYc = min(a.rpc, NY - ck.row0);
  ck.NP = (ck.NYc + 1) >> 1;
  ck.NS = ck.NP * ck.NX;
  ck.iters = (ck.NS + 31) >> 5;
  ck.cxlo = cx - Rx;
  ck.cylo = cy - Ry;

  uint32_t *s_org = (uint32_t *)smem;       // [64][16 * PB] words
  uint32_t *s_ref = s_org + 64 * 16 * PB;   // [nrows][rsw][PXW pixel phases]
  uint32_t *s_x = s_ref + a.csz;       // 2 x [5][16][64]
  const int nrows = ck.NYc + 63;
  // u16 only: negated 4x4 box sums of the window [nrows - 3][rsw * PXW], summed-area table of org's 4x4 cells
  uint32_t *s_rcol = s_x + 2 * 5 * 16 * 64;  // [256] rate of the column part of a candidate: pmv0 | (pmv1 + 1) << 16
  uint32_t *s_rrow = s_rcol + 256;          // [rpc] row part
  uint32_t *s_sat = s_rrow + a.rpc;         // [17][17]
  int16_t *s_nbox = (int16_t *)(s_sat + 17 * 17 + 3);

  // ---- stage the org region and phase 0 of the window (clamped into the allocation:
  //      pixels outside it never belong to a legal candidate) ---------------------------
  {
    const uint8_t *po = (const uint8_t *)a.org.p0, *pr = (const uint8_t *)a.ref.p0;
    const int oxmin = -a.org.xorigin, oxmax = a.org.alloc_w - a.org.xorigin - 1;
    const int oymin = -a.org.yorigin, oymax = a.org.alloc_h - a.org.yorigin - 1;
    for (int i = tid; i < 64 * 64; i += T2) {
      const int y = min(max(job.po_y + (i >> 6), oymin), oymax), x = min(max(job.po_x + (i & 63), oxmin), oxmax);
      if (PB == 1) ((uint8_t *)s_org)[i] = po[(long long)y * a.org.stride + x];
      else ((uint16_t *)s_org)[i] = ((const uint16_t *)po)[(long long)y * a.org.stride + x];
    }
    const int rxmin = -a.ref.xorigin, rxmax = a.ref.alloc_w - a.ref.xorigin - 1;
    const int rymin = -a.ref.yorigin, rymax = a.ref.alloc_h - a.ref.yorigin - 1;
    const int x0 = job.po_x + ck.cxlo, y0 = job.po_y + ck.cylo + ck.row0;
    // one thread = one word (4 / PB pixels) of one row
    for (int wy = q; wy < nrows; wy += T2 / 32) {
      const int y = min(max(y0 + wy, rymin), rymax);
      const uint8_t *row = pr + (long long)y * a.ref.stride * PB;
      for (int w = lane; w < rsw; w += 32) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < PXW; b++) {
          const int x = min(max(x0 + PXW * w + b, rxmin), rxmax);
          v |= (PB == 1 ? (uint32_t)row[x] : (uint32_t)((const uint16_t *)row)[x]) << (8 * PB * b);
        }
        s_ref[(wy * rsw + w) * PXW] = v;
      }
    }
  }
  __syncthreads();
  // ---- the other phases: the same row shifted left by 1..PXW-1 pixels --------------------
  for (int wy = q; wy < nrows; wy += T2 / 32)
    for (int w = lane; w < rsw; w += 32) {
      const int i = wy * rsw + w;
      const uint32_t lo = s_ref[i * PXW], hi = (w + 1 < rsw) ? s_ref[(i + 1) * PXW] : 0u;
#pragma unroll
      for (int ph = 1; ph < PXW; ph++) s_ref[i * PXW + ph] = __funnelshift_r(lo, hi, 8 * PB * ph);
    }
  {  // get_mv_rate (src/me.rs:1005-1021) per column and per row of the chunk's candidates
    const int hp = a.p.allow_high_precision_mv;
    for (int i = tid; i < ck.NX; i += T2) {
      const int mvc = 8 * (ck.cxlo + i);
      s_rcol[i] = rate1(mvc - job.pmv[0].col, hp) | ((rate1(mvc - job.pmv[1].col, hp) + 1u) << 16);
    }
    for (int i = tid; i < ck.NYc; i += T2) {
      const int mvr = 8 * (ck.cylo + ck.row0 + i);
      s_rrow[i] = rate1(mvr - job.pmv[0].row, hp) | (rate1(mvr - job.pmv[1].row, hp) << 16);
    }
  }
  // sums of org's 4x4 cells: up to 10 bits for the block sums of chunk 0 (-> s_x, free until the search starts);
  // at 12 bits every chunk takes them off its cells (-> s_sat, which has no other use there)
  if (PB == 2 && (D12 || chunk == 0)) {
    if (tid < 256) {
      const uint16_t *o = (const uint16_t *)s_org + (tid >> 4) * 4 * 64 + (tid & 15) * 4;
      uint32_t sum = 0;
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) sum += o[r * 64 + c];
      (D12 ? s_sat : s_x)[tid] = sum;
    }
  }
  __syncthreads();
  if (PB == 2) {
    // box(y, x) = sum of window pixels [y, y + 4) x [x, x + 4): a thread owns a column and a segment of rows and
    // slides down it with the row sums h4 of the last four rows in a ring
    const int bw = rsw * PXW, nby = nrows - 3;
    const int nseg = max(1, T2 / bw), seg_rows = (nby + nseg - 1) / nseg;
    for (int t = tid; t < bw * nseg; t += T2) {
      const int sg = t / bw, x = t - sg * bw;
      const int ylo = sg * seg_rows, yhi = min(nby, ylo + seg_rows);
      if (x > bw - 4) continue;
      const uint32_t *col = s_ref + (x >> 1) * PXW + (x & 1);
      auto h4 = [&](int y) {
        const uint32_t v = col[y * rsw * PXW] + col[(y * rsw + 1) * PXW];  // two pixel pairs, halves stay apart
        return (int)((v & 0xffffu) + (v >> 16));
      };
      if (ylo < yhi) {
        int h0 = h4(ylo), h1 = h4(ylo + 1), h2 = h4(ylo + 2);
        for (int y = ylo; y < yhi; y++) {
          const int h3 = h4(y + 3);
          // (12 bits: a box reaches 16 * 4095 = 65520, kept as 32760 - box; the offset leaves with sum org, see sa[][])
          s_nbox[y * bw + x] = (int16_t)((D12 ? 32760 : 0) - (h0 + h1 + h2 + h3));
          h0 = h1;
          h1 = h2;
          h2 = h3;
        }
      }
    }
    // summed-area table of the cell sums: sat[Y][X] = sum of cells [0, Y) x [0, X); the minima are taken on
    // 256 * (SAD + sum org of the block) + rl, and the block sums go to sb2_finalize_kernel once per superblock
    if (!D12 && chunk == 0) {
      if (tid < 17 * 17) {
        const int Y = tid / 17, X = tid - Y * 17;
        uint32_t sum = 0;
        for (int y = 0; y < Y; y++)
          for (int x = 0; x < X; x++) sum += s_x[y * 16 + x];
        s_sat[tid] = sum;
      }
      __syncthreads();
      uint32_t *bias = a.bias + ji * NSLOT;
      for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
        const int cnt = sb2_count(b);
        if (cnt == 0) continue;
        const int cw = blk_w_of(b) >> 2, ch = blk_h_of(b) >> 2, per_row = 16 / cw, off = sb2_offset(b);
        for (int k = tid; k < cnt; k += T2) {
          const int y0 = (k / per_row) * ch, x0 = (k % per_row) * cw, y1 = y0 + ch, x1 = x0 + cw;
          bias[off + k] = s_sat[y1 * 17 + x1] - s_sat[y0 * 17 + x1] - s_sat[y1 * 17 + x0] + s_sat[y0 * 17 + x0];
        }
      }
    }
    __syncthreads();
  }

  Geo2 g;
  g.RX = job.po_x;
  g.RY = job.po_y;
  g.QX = job.po_x + (q & 3) * 16;
  g.QY = job.po_y + (q >> 2) * 16;
  g.W4 = a.p.w_in_b * 4;
  g.H4 = a.p.h_in_b * 4;
  // every candidate of the chunk is legal for every block of the region (the most
  // restrictive blocks are the 4x4s on the region's edges)
  const int oylo = ck.cylo + ck.row0, oyhi = oylo + ck.NYc - 1;
  const bool fast = (job.po_x + 60 + ck.cxlo + 4 >= -16) && (job.po_x + cx + Rx <= g.W4 + 16) &&
                    (job.po_x + ck.cxlo + 4 >= -16) && (job.po_x + 60 + cx + Rx <= g.W4 + 16) &&
                    (job.po_y + oylo + 4 >= -16) && (job.po_y + 60 + oyhi <= g.H4 + 16) &&
                    (job.po_y + 60 + oylo + 4 >= -16) && (job.po_y + oyhi <= g.H4 + 16);

  uint32_t best[44], pc[5], ps[5], bc[5], bs[5];
#pragma unroll
  for (int i = 0; i < 44; i++) best[i] = NONE32;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    pc[i] = bc[i] = NONE32;
    ps[i] = bs[i] = 0;
  }
  if (fast) sb2_search<false, RSW, PB, D12>(a, job, ck, s_org, s_ref, s_x, s_nbox, s_sat, s_rcol, s_rrow, best, pc, ps, bc, bs, g);
  else sb2_search<true, RSW, PB, D12>(a, job, ck, s_org, s_ref, s_x, s_nbox, s_sat, s_rcol, s_rrow, best, pc, ps, bc, bs, g);

  // ---- reduce over the 32 lanes and write the partial keys ------------------------------
  unsigned long long *out = a.partial + wi * NSLOT;
  const int qx16 = (q & 3) * 16, qy16 = (q >> 2) * 16;
  const uint32_t inv_m2 = (65536u + 2 * a.M - 1) / (2 * a.M);
#pragma unroll
  for (int bi = 0; bi < 49; bi++) {
    uint32_t cost, seq;
    if (bi < 44) {
      cost = best[bi] == NONE32 ? NONE32 : (best[bi] >> SEQ_BITS);
      seq = best[bi] & ((1u << SEQ_BITS) - 1);
    } else {
      cost = pc[bi - 44];
      seq = ps[bi - 44];
    }
    const unsigned long long key = warp_best(cost, decode_ridx(ck, seq, lane, a.M, inv_m2));
    if (lane == 0) {
      int bsize, lx, ly;
      if (bi < 16) { bsize = R1C_BLOCK_4X4; lx = 4 * (bi & 3); ly = 4 * (bi >> 2); }
      else if (bi < 24) { bsize = R1C_BLOCK_4X8; lx = 4 * ((bi - 16) & 3); ly = 8 * ((bi - 16) >> 2); }
      else if (bi < 32) { bsize = R1C_BLOCK_8X4; lx = 8 * ((bi - 24) & 1); ly = 4 * ((bi - 24) >> 1); }
      else if (bi < 36) { bsize = R1C_BLOCK_8X8; lx = 8 * ((bi - 32) & 1); ly = 8 * ((bi - 32) >> 1); }
      else if (bi < 40) { bsize = R1C_BLOCK_4X16; lx = 4 * (bi - 36); ly = 0; }
      else if (bi < 44) { bsize = R1C_BLOCK_16X4; lx = 0; ly = 4 * (bi - 40); }
      else if (bi < 46) { bsize = R1C_BLOCK_8X16; lx = 8 * (bi - 44); ly = 0; }
      else if (bi < 48) { bsize = R1C_BLOCK_16X8; lx = 0; ly = 8 * (bi - 46); }
      else { bsize = R1C_BLOCK_16X16; lx = 0; ly = 0; }
      const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
      out[sb2_offset(bsize) + ((qy16 + ly) / bh) * (64 / bw) + (qx16 + lx) / bw] = key;
    }
  }
#pragma unroll
  for (int s = 0; s < 5; s++) {
    const BigSlot b = big_slot(q, s);
    if (b.bsize >= 0) {  // warp-uniform
      const unsigned long long key = warp_best(bc[s], decode_ridx(ck, bs[s], lane, a.M, inv_m2));
      if (lane == 0) {
        const int bw = blk_w_of(b.bsize), bh = blk_h_of(b.bsize);
        out[sb2_offset(b.bsize) + (b.ly / bh) * (64 / bw) + b.lx / bw] = key;
      }
    }
  }
}

// min over the chunks -> (mv, cost); entries of blocks that do not exist in the frame or
// are not in the job's mask are left untouched
struct FinArgs {
  const unsigned long long *partial;
  const uint32_t *bias;  // u16 planes: sum of org per block, taken off the cost (x 256); else nullptr
  const r1c_sb_job *jobs;
  r1c_me_result *results;
  const int2 *map;  // optional: (source slot, destination index)
  long long n_out;
  int split, NX, Rx, Ry, W4, H4;
};
__global__ void sb2_finalize_kernel(FinArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_out) return;
  const long long src = a.map ? a.map[i].x : i;
  const long long dst = a.map ? a.map[i].y : i;
  const long long ji = src / NSLOT;
  const int slot = (int)(src - ji * NSLOT);
  const r1c_sb_job job = a.jobs[ji];
  int bsize = 0, off = 0;
  for (int b = 0; b < R1C_BLOCK_SIZES_ALL; b++) {
    const int c = sb2_count(b);
    if (slot < off + c) {
      bsize = b;
      break;
    }
    off += c;
  }
  if (!((job.bsize_mask >> bsize) & 1u)) return;
  const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
  const int k = slot - off, rx = (k % (64 / bw)) * bw, ry = (k / (64 / bw)) * bh;
  if (job.po_x + rx + bw > a.W4 || job.po_y + ry + bh > a.H4) return;
  unsigned long long key = ~0ull;
  for (int c = 0; c < a.split; c++) {
    const unsigned long long v = a.partial[(ji * a.split + c) * NSLOT + slot];
    key = v < key ? v : key;
  }
  r1c_me_result r;
  r.reserved = 0;
  if (key == ~0ull) {
    r.mv.row = 0;
    r.mv.col = 0;
    r.cost = ~0ull;
  } else {
    const uint32_t ridx = (uint32_t)key;
    const int dy = (int)(ridx / (uint32_t)a.NX), dx = (int)(ridx - (uint32_t)dy * a.NX);
    r.mv.row = (int16_t)(8 * (job.cmv.row / 8 - a.Ry + dy));
    r.mv.col = (int16_t)(8 * (job.cmv.col / 8 - a.Rx + dx));
    r.cost = (key >> 32) - (a.bias ? 256ull * a.bias[ji * NSLOT + slot] : 0ull);
  }
  a.results[dst] = r;
}

}  // namespace

// layout parameters of the kernel for a search range; false when this kernel cannot
// take the job (the caller falls back to me_sb.cu)
struct Sb2Plan {
  int split, rpc, rsw, csz, M;
  size_t smem;
};
typedef void (*sb2_kernel_fn)(Sb2Args);
static sb2_kernel_fn sb2_kernel_of(int px_bytes, int bit_depth, int rsw) {
  if (px_bytes == 1) return rsw == 48 ? me_sb2_kernel<48, 1> : me_sb2_kernel<0, 1>;
  if (bit_depth > 10) return rsw == 96 ? me_sb2_kernel<96, 2, true> : me_sb2_kernel<0, 2, true>;
  return rsw == 96 ? me_sb2_kernel<96, 2> : me_sb2_kernel<0, 2>;
}
// the block sums of org that leave the cost in sb2_finalize_kernel: u16 planes up to 10 bits only
static size_t sb2_bias_bytes(int px_bytes, int bit_depth, int64_t n) {
  return px_bytes == 2 && bit_depth <= 10 ? (size_t)n * NSLOT * sizeof(uint32_t) : 0;
}
static bool sb2_plan(const r1c_ctx *ctx, const r1c_sb_params *p, int px_bytes, int bit_depth, int64_t n, Sb2Plan *pl) {
  // u16 planes up to 10 bits: packed-key SADs (8x8 <= 64 * 1023) and SAD * 256 + rl of a 64x64 block fit 32 bits;
  // 12 bits: the D12 variant (true SADs in the keys)
  if ((px_bytes != 1 && px_bytes != 2) || p->step != 1) return false;
  const bool d12 = px_bytes == 2 && bit_depth > 10;
  if (p->lambda > (d12 ? 16000u : 60000u)) return false;  // rl >> 8 below 2^14 (rate <= 65); D12: 256 * SAD(64x64) + rl < 2^32 - 1
  const int seq_bits = px_bytes == 1 ? Px<1>::SEQ_BITS : (d12 ? Px<2, true>::SEQ_BITS : Px<2>::SEQ_BITS), pxw = 4 / px_bytes;
  const int NX = 2 * p->range_x + 1, NY = 2 * p->range_y + 1;
  if (p->range_y < 1 || NX > 255) return false;
  const int M = (NX + 31) / 32;
  const int winW = 64 + 2 * p->range_x;
  const int rsw = (winW + pxw - 1) / pxw;  // a copy row holds pixels [p, p + pxw * rsw)
  const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
  for (int split = 1; split <= 32; split++) {
    const int rpc = 2 * ((NY + 2 * split - 1) / (2 * split));
    const int chunks = (NY + rpc - 1) / rpc;
    if (chunks != split) continue;
    const int last = NY - (split - 1) * rpc;
    if (last < 2) continue;
    const int NP = rpc / 2;
    if (NP * 2 * M > (1 << seq_bits)) continue;  // seq field
    if ((long long)(last + 1) / 2 * NX < 32) continue;  // every lane needs a slot of its own
    const int nrows = rpc + 63;
    const int csz = nrows * rsw * pxw;
    size_t smem = (size_t)64 * 64 * px_bytes + (size_t)csz * 4 + (size_t)2 * 5 * 16 * 64 * 4 + (size_t)(256 + rpc) * 4;
    if (px_bytes == 2) smem += (size_t)(17 * 17 + 3) * 4 + (size_t)(nrows - 3) * rsw * pxw * 2;  // s_sat, s_nbox
    if (smem > 220 * 1024) continue;
    // enough work items to balance the SMs: stop once the tail is small
    pl->split = split;
    pl->rpc = rpc;
    pl->rsw = rsw;
    pl->csz = csz;
    pl->M = M;
    pl->smem = smem;
    const long long items = n * split;
    const long long waves = (items + sms - 1) / sms;
    if ((double)items / (double)(waves * sms) > 0.93 || items >= 8ll * sms) return true;
  }
  return pl->split > 0 && pl->smem > 0;
}

// search + finalize.  d_map == nullptr: results is the [n][849] layout.
int r1c_sb2_search(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_sb_params *p, int bit_depth,
                   const r1c_sb_job *d_jobs, int64_t n, r1c_me_result *d_results, const int2 *d_map, int64_t n_map,
                   int *handled) {
  *handled = 0;
  Sb2Plan pl = {};
  const char *force = getenv("R1C_ME_SB_V1");
  if ((force && force[0] == '1') || org->px_bytes != ref->px_bytes || !sb2_plan(ctx, p, org->px_bytes, bit_depth, n, &pl))
    return R1C_OK;
  *handled = 1;
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *d_partial;
  const size_t partial_bytes = (size_t)n * pl.split * NSLOT * sizeof(unsigned long long);
  const size_t bias_bytes = sb2_bias_bytes(org->px_bytes, bit_depth, n);
  int rc = r1c_scratch_dev(ctx, 15, partial_bytes + bias_bytes, &d_partial);
  if (rc) return rc;
  Sb2Args a;
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.p = *p;
  a.jobs = d_jobs;
  a.partial = (unsigned long long *)d_partial;
  a.bias = bias_bytes ? (uint32_t *)((uint8_t *)d_partial + partial_bytes) : nullptr;
  a.n = n;
  a.split = pl.split;
  a.rpc = pl.rpc;
  a.rsw = pl.rsw;
  a.csz = pl.csz;
  a.M = pl.M;
  // 48 / 96 words per window row: the +-64 px window
  auto kern = sb2_kernel_of(org->px_bytes, bit_depth, pl.rsw);
  R1C_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
  r1c_prof_begin(ctx, R1C_PROF_ME_SB);
  kern<<<(unsigned)(n * pl.split), T2, pl.smem, ctx->stream>>>(a);
  r1c_prof_end(ctx, R1C_PROF_ME_SB);
  R1C_LAUNCH_CHECK(ctx);
  FinArgs fa;
  fa.partial = a.partial;
  fa.bias = a.bias;
  fa.jobs = d_jobs;
  fa.results = d_results;
  fa.map = d_map;
  fa.n_out = d_map ? n_map : n * NSLOT;
  fa.split = pl.split;
  fa.NX = 2 * p->range_x + 1;
  fa.Rx = p->range_x;
  fa.Ry = p->range_y;
  fa.W4 = p->w_in_b * 4;
  fa.H4 = p->h_in_b * 4;
  if (fa.n_out > 0) {
    sb2_finalize_kernel<<<(unsigned)((fa.n_out + 255) / 256), 256, 0, ctx->stream>>>(fa);
    R1C_LAUNCH_CHECK(ctx);
  }
  return R1C_OK;
}

// ---------------------------------------------------------------------------------------------
// r1c_me_sb_batch_dev_noalloc -- the search above for callers that may not allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The partial keys (and the org sums of u16
// planes) live in ONE file-scope device array of R1C_SB_NOALLOC_SCRATCH_BYTES instead of the
// context's grow-on-demand scratch: every GPU that loads the library holds it until exit, every
// call on a device shares it, and calls that overlap on the GPU (two streams, two contexts)
// race on it.  Every slot the finalize kernel reads is written by the search kernel of the same
// call first (849 keys per work item, 849 sums per job), so nothing has to start at zero and
// the contents left by the previous call are never read.
// In host code `g_sb2_noalloc_scratch` is only a stand-in for the symbol; the device address
// comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_sb2_noalloc_scratch[R1C_SB_NOALLOC_SCRATCH_BYTES];

#include <mutex>
static std::mutex g_noalloc_mu;
static void *g_noalloc_ptr[64];  // per device ordinal; nullptr = not looked up yet

extern "C" int r1c_me_sb_batch_dev_noalloc(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                           const r1c_sb_params *p, const r1c_sb_job *d_jobs, int64_t n,
                                           int bit_depth, r1c_me_result *d_results) {
  if (!ctx || !org || !ref || !p) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes, R1C_ERR_ARG, "org/ref pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth,
              org->px_bytes);
  R1C_REQUIRE(ctx, p->range_x >= 0 && p->range_y >= 0 && p->w_in_b > 0 && p->h_in_b > 0, R1C_ERR_ARG, "bad range / frame size");
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 31), R1C_ERR_ARG, "bad job count");
  R1C_REQUIRE(ctx, n == 0 || (d_jobs && d_results), R1C_ERR_ARG, "null job/result array");
  R1C_REQUIRE(ctx, ctx->device >= 0 && ctx->device < 64, R1C_ERR_ARG, "device ordinal %d not supported by the no-alloc path", ctx->device);
  Sb2Plan pl = {};
  R1C_REQUIRE(ctx, sb2_plan(ctx, p, org->px_bytes, bit_depth, n, &pl), R1C_ERR_UNIMPLEMENTED,
              "no-alloc search: step 1, lambda <= 60000 (16000 at 12 bits) and a window this kernel's layout can take only "
              "(range %dx%d, step %d, lambda %u)", p->range_x, p->range_y, p->step, p->lambda);
  if (n == 0) return R1C_OK;
  const size_t partial_bytes = (size_t)n * pl.split * NSLOT * sizeof(unsigned long long);
  const size_t bias_bytes = sb2_bias_bytes(org->px_bytes, bit_depth, n);
  R1C_REQUIRE(ctx, partial_bytes + bias_bytes <= (size_t)R1C_SB_NOALLOC_SCRATCH_BYTES, R1C_ERR_NOMEM,
              "no-alloc search: %lld superblocks x %d chunks need %zu bytes of scratch, the static array holds %u "
              "(R1C_SB_NOALLOC_SCRATCH_BYTES); split the job list or use r1c_me_sb_batch_dev",
              (long long)n, pl.split, partial_bytes + bias_bytes, (unsigned)R1C_SB_NOALLOC_SCRATCH_BYTES);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  void *d_scratch;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
      void *sym = nullptr;
      R1C_CUDA(ctx, cudaGetSymbolAddress(&sym, g_sb2_noalloc_scratch));
      R1C_REQUIRE(ctx, sym != nullptr, R1C_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
      g_noalloc_ptr[ctx->device] = sym;
    }
    d_scratch = g_noalloc_ptr[ctx->device];
  }
  Sb2Args a;
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.p = *p;
  a.jobs = d_jobs;
  a.partial = (unsigned long long *)d_scratch;
  a.bias = bias_bytes ? (uint32_t *)((uint8_t *)d_scratch + partial_bytes) : nullptr;
  a.n = n;
  a.split = pl.split;
  a.rpc = pl.rpc;
  a.rsw = pl.rsw;
  a.csz = pl.csz;
  a.M = pl.M;
  auto kern = sb2_kernel_of(org->px_bytes, bit_depth, pl.rsw);
  R1C_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
  kern<<<(unsigned)(n * pl.split), T2, pl.smem, ctx->stream>>>(a);
  R1C_LAUNCH_CHECK(ctx);
  FinArgs fa;
  fa.partial = a.partial;
  fa.bias = a.bias;
  fa.jobs = d_jobs;
  fa.results = d_results;
  fa.map = nullptr;
  fa.n_out = n * NSLOT;
  fa.split = pl.split;
  fa.NX = 2 * p->range_x + 1;
  fa.Rx = p->range_x;
  fa.Ry = p->range_y;
  fa.W4 = p->w_in_b * 4;
  fa.H4 = p->h_in_b * 4;
  sb2_finalize_kernel<<<(unsigned)((fa.n_out + 255) / 256), 256, 0, ctx->stream>>>(fa);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}
