// me_sb.cu -- whole-superblock full search.
//
// What it computes: for a 64x64 region of the frame, the result of the
// reference's full_search (src/me.rs:943-990) for EVERY block of every size
// 4x4 .. 64x64 (19 BlockSize shapes, 849 blocks) against the windows that
// FullSearch::full_pixel_me builds (src/me.rs:533-546) -- in one pass over the
// reference window.  A 4x4 SAD is computed once per (cell, candidate) and
// summed up the block-size hierarchy, instead of once per block size.
//
// Mapping (one CTA of 256 threads per superblock):
//   thread = (16x16 quadrant q of the region, candidate lane c of 16).
//   Per iteration the 16 lanes take 16 consecutive candidates of the raster
//   scan.  A thread computes the 16 cell SADs of its quadrant for its candidate
//   with VABSDIFF4.U8.ACC on words realigned by a funnel shift, sums them to
//   the 49 blocks that lie inside a quadrant and keeps a running best for each
//   of them in registers.  The five partial sums every larger block is made of
//   (16x16, 8x16 left/right, 16x8 top/bottom) go through a double-buffered
//   shared-memory exchange; the 65 blocks that span quadrants are handled by
//   fixed (block, lane) pairs.
//   Running best = one 64-bit key  cost << 20 | iteration  (IMAD.WIDE builds
//   it from the SAD), so min() on the key keeps the FIRST minimum of the raster
//   scan exactly like the reference's strict `<`.
// Shared memory: the org region (4/8 KB), the reference window
// (64+2Rx) x (64+2Ry) pixels, the exchange buffers (10 KB).
// INT-bound: the planes are read once per superblock from L2/HBM; the work is
// 2 ops per (pixel, candidate) at the 4x4 layer.
#include "me_common.cuh"

namespace {

constexpr int SB_THREADS = 256;
constexpr unsigned long long KEY_MAX = ~0ull;
constexpr int IT_BITS = 20;

struct SbArgs {
  PlaneView org, ref;
  r1c_sb_params p;
  const r1c_sb_job *jobs;
  r1c_me_result *results;
  long long n;
  int rstride;  // bytes per row of the staged reference window
};

// offsets of each BlockSize inside the 849-entry per-superblock result array
// (BlockSize order; sizes above 64x64 have none)
__host__ __device__ inline int sb_count(int b) {
  const int w = blk_w_of(b), h = blk_h_of(b);
  return (w > 64 || h > 64) ? 0 : (64 / w) * (64 / h);
}
__host__ __device__ inline int sb_offset(int b) {
  int o = 0;
  for (int i = 0; i < b; i++) o += sb_count(i);
  return o;
}

// the 65 blocks that span more than one 16x16 quadrant
struct BigBlk {
  unsigned char bsize, qx0, qy0, nqx, nqy, part, lx, ly;  // lx, ly in units of 4 px
};
__constant__ BigBlk c_big[65];

__device__ __forceinline__ uint32_t rate_fast(int16_t diff, int allow_hp) {
  const int d = allow_hp ? (int)diff : ((int)diff >> 1);
  if (d == 0) return 0;
  const uint32_t a = (uint32_t)(d < 0 ? -d : d) & 0xffffu;
  return 2u * (uint32_t)(32 - __clz(a));
}

struct Geo {
  int QX, QY;      // absolute position of the thread's quadrant
  int W4, H4;      // frame size in pixels (w_in_b * 4)
  int cxlo, cylo;  // first offset of the nominal window: cx - Rx, cy - Ry
  int step;
};

// candidate validity for a block at quadrant-local (lx, ly) of size (bw, bh) and
// MV offset (ox, oy): the window of FullSearch::full_pixel_me after get_mv_range
//   -(bx + 16 + bw) <= ox <= W4 - bx + 16      (same in y)
// and, for step 2, the lattice anchored at the block's own x_lo / y_lo.
__device__ __forceinline__ bool cand_valid(const Geo &g, int lx, int ly, int bw, int bh, int ox, int oy) {
  const int bx = g.QX + lx, by = g.QY + ly;
  bool v = (bx + ox + bw >= -16) && (bx + ox <= g.W4 + 16) && (by + oy + bh >= -16) && (by + oy <= g.H4 + 16);
  if (g.step == 2) {
    const int xlo = max(g.cxlo, -(bx + 16 + bw)), ylo = max(g.cylo, -(by + 16 + bh));
    v = v && (((ox - xlo) & 1) == 0) && (((oy - ylo) & 1) == 0);
  }
  return v;
}

template <int PB, bool FAST>
__device__ __forceinline__ void sb_search(const SbArgs &a, const r1c_sb_job &job, const uint32_t *s_org,
                                          const uint8_t *s_ref, uint32_t *s_x, unsigned long long (&best)[49],
                                          unsigned long long (&bbest)[5], const Geo &g, int NX, int NC, int q, int c) {
  constexpr int WPQ = 16 * PB / 4;  // words per quadrant row
  constexpr int OSTR = 64 * PB / 4; // org words per row
  const int qx = q & 3, qy = q >> 2;
  const int tid = threadIdx.x;
  const int iters = (NC + 15) >> 4;
  for (int it = 0; it < iters; it++) {
    const int ci = it * 16 + c;
    const bool live = ci < NC;
    const int dy = live ? ci / NX : 0, dx = live ? ci - dy * NX : 0;
    const int ox = g.cxlo + dx, oy = g.cylo + dy;
    // ---- 16 cell SADs of the quadrant --------------------------------------
    uint32_t cell[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) cell[i][k] = 0;
    {
      const int bxo = (qx * 16 + dx) * PB;  // byte offset inside a window row
      const uint32_t sh = (uint32_t)(bxo & 3) * 8;
      const uint32_t *rrow = (const uint32_t *)(s_ref + (size_t)(qy * 16 + dy) * a.rstride) + (bxo >> 2);
      const uint32_t *orow = s_org + (qy * 16) * OSTR + qx * WPQ;
      const int rstep = a.rstride >> 2;
#pragma unroll
      for (int r = 0; r < 16; r++) {
        uint32_t w[WPQ + 1];
#pragma unroll
        for (int k = 0; k < WPQ + 1; k++) w[k] = rrow[r * rstep + k];
        const uint4 o0 = *(const uint4 *)(orow + r * OSTR);
        uint32_t ow[WPQ];
        ow[0] = o0.x; ow[1] = o0.y; ow[2] = o0.z; ow[3] = o0.w;
        if (PB == 2) {
          const uint4 o1 = *(const uint4 *)(orow + r * OSTR + 4);
          ow[WPQ - 4] = o1.x; ow[WPQ - 3] = o1.y; ow[WPQ - 2] = o1.z; ow[WPQ - 1] = o1.w;
        }
#pragma unroll
        for (int k = 0; k < WPQ; k++) {
          const uint32_t rv = __funnelshift_r(w[k], w[k + 1], sh);
          cell[r >> 2][k * 4 / WPQ] = me::sad_word<PB>(ow[k], rv, cell[r >> 2][k * 4 / WPQ]);
        }
      }
    }
    // ---- rate term and key base ----------------------------------------------
    r1c_mv mv;
    mv.row = (int16_t)(8 * oy);
    mv.col = (int16_t)(8 * ox);
    const uint32_t r1 = rate_fast((int16_t)(mv.row - job.pmv[0].row), a.p.allow_high_precision_mv) +
                        rate_fast((int16_t)(mv.col - job.pmv[0].col), a.p.allow_high_precision_mv);
    const uint32_t r2 = rate_fast((int16_t)(mv.row - job.pmv[1].row), a.p.allow_high_precision_mv) +
                        rate_fast((int16_t)(mv.col - job.pmv[1].col), a.p.allow_high_precision_mv) + 1;
    const unsigned long long rc = (unsigned long long)min(r1, r2) * a.p.lambda;
    const unsigned long long rck = live ? ((rc << IT_BITS) | (unsigned long long)it) : KEY_MAX;

#define SB_UPD(BI, SAD, LX, LY, BW, BH)                                                        \
  do {                                                                                         \
    if (live && (FAST || cand_valid(g, LX, LY, BW, BH, ox, oy))) {                             \
      const unsigned long long key_ = (unsigned long long)(SAD) * (1ull << (8 + IT_BITS)) + rck; \
      if (key_ < best[BI]) best[BI] = key_;                                                    \
    }                                                                                          \
  } while (0)

    // ---- the 49 blocks inside the quadrant -----------------------------------
    // index layout: 4x4 [0,16) 4x8 [16,24) 8x4 [24,32) 8x8 [32,36) 8x16 [36,38)
    //               16x8 [38,40) 16x16 [40] 4x16 [41,45) 16x4 [45,49)
    uint32_t v48[2][4], h84[4][2], s88[2][2];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) SB_UPD(i * 4 + k, cell[i][k], 4 * k, 4 * i, 4, 4);
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int k = 0; k < 4; k++) {
        v48[i][k] = cell[2 * i][k] + cell[2 * i + 1][k];
        SB_UPD(16 + i * 4 + k, v48[i][k], 4 * k, 8 * i, 4, 8);
      }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int k = 0; k < 2; k++) {
        h84[i][k] = cell[i][2 * k] + cell[i][2 * k + 1];
        SB_UPD(24 + i * 2 + k, h84[i][k], 8 * k, 4 * i, 8, 4);
      }
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int k = 0; k < 2; k++) {
        s88[i][k] = v48[i][2 * k] + v48[i][2 * k + 1];
        SB_UPD(32 + i * 2 + k, s88[i][k], 8 * k, 8 * i, 8, 8);
      }
    const uint32_t c816l = s88[0][0] + s88[1][0], c816r = s88[0][1] + s88[1][1];  // 8 wide, 16 high
    const uint32_t c168t = s88[0][0] + s88[0][1], c168b = s88[1][0] + s88[1][1];  // 16 wide, 8 high
    const uint32_t c1616 = c816l + c816r;
    SB_UPD(36, c816l, 0, 0, 8, 16);
    SB_UPD(37, c816r, 8, 0, 8, 16);
    SB_UPD(38, c168t, 0, 0, 16, 8);
    SB_UPD(39, c168b, 0, 8, 16, 8);
    SB_UPD(40, c1616, 0, 0, 16, 16);
#pragma unroll
    for (int k = 0; k < 4; k++) SB_UPD(41 + k, v48[0][k] + v48[1][k], 4 * k, 0, 4, 16);
#pragma unroll
    for (int i = 0; i < 4; i++) SB_UPD(45 + i, h84[i][0] + h84[i][1], 0, 4 * i, 16, 4);

    // ---- exchange of the quadrant partial sums ---------------------------------
    uint32_t *xb = s_x + (it & 1) * (16 * 5 * 16);
    xb[(q * 5 + 0) * 16 + c] = c1616;
    xb[(q * 5 + 1) * 16 + c] = c816l;
    xb[(q * 5 + 2) * 16 + c] = c816r;
    xb[(q * 5 + 3) * 16 + c] = c168t;
    xb[(q * 5 + 4) * 16 + c] = c168b;
    __syncthreads();
    // ---- the 65 blocks spanning quadrants: pair p = (block p>>4, lane p&15) ----
#pragma unroll
    for (int rd = 0; rd < 5; rd++) {
      const int p = tid + rd * SB_THREADS;
      if (p < 65 * 16) {
        const BigBlk bb = c_big[p >> 4];
        const int lc = p & 15;  // == c (256 is a multiple of 16)
        uint32_t s = 0;
        for (int yy = 0; yy < bb.nqy; yy++)
          for (int xx = 0; xx < bb.nqx; xx++) s += xb[(((bb.qy0 + yy) * 4 + bb.qx0 + xx) * 5 + bb.part) * 16 + lc];
        // this pair's candidate is lane lc of the same iteration: same (ox, oy, rck) as
        // the thread's own lane because lc == c
        bool ok = live;
        if (!FAST && ok) {
          Geo g0 = g;
          g0.QX = g.QX - qx * 16;  // region origin
          g0.QY = g.QY - qy * 16;
          ok = cand_valid(g0, bb.lx * 4, bb.ly * 4, blk_w_of(bb.bsize), blk_h_of(bb.bsize), ox, oy);
        }
        if (ok) {
          const unsigned long long key_ = (unsigned long long)s * (1ull << (8 + IT_BITS)) + rck;
          if (key_ < bbest[rd]) bbest[rd] = key_;
        }
      }
    }
  }
#undef SB_UPD
}

// lexicographic min over the 16 lanes of a half-warp on (key, lane)
__device__ __forceinline__ void reduce16(unsigned long long &key, int &lane) {
#pragma unroll
  for (int m = 1; m < 16; m <<= 1) {
    const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, m);
    const int ol = __shfl_xor_sync(0xffffffffu, lane, m);
    if (ok < key || (ok == key && ol < lane)) {
      key = ok;
      lane = ol;
    }
  }
}

__device__ __forceinline__ void write_result(const SbArgs &a, const r1c_sb_job &job, long long ji, int bsize,
                                             int rx, int ry, unsigned long long key, int lane, int NX,
                                             const Geo &g) {
  // rx, ry: block origin relative to the region
  const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
  if (!((job.bsize_mask >> bsize) & 1u)) return;
  if (job.po_x + rx + bw > g.W4 || job.po_y + ry + bh > g.H4) return;  // block does not exist in this frame
  r1c_me_result r;
  r.reserved = 0;
  if (key == KEY_MAX) {
    r.mv.row = 0;
    r.mv.col = 0;
    r.cost = ~0ull;
  } else {
    const int it = (int)(key & ((1u << IT_BITS) - 1));
    const int ci = it * 16 + lane;
    const int dy = ci / NX, dx = ci - dy * NX;
    r.mv.row = (int16_t)(8 * (g.cylo + dy));
    r.mv.col = (int16_t)(8 * (g.cxlo + dx));
    r.cost = key >> IT_BITS;
  }
  const int idx = (ry / bh) * (64 / bw) + rx / bw;
  a.results[ji * R1C_SB_RESULTS + sb_offset(bsize) + idx] = r;
}

template <int PB>
__global__ void __launch_bounds__(SB_THREADS, 1) me_sb_kernel(SbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  const long long ji = blockIdx.x;
  const r1c_sb_job job = a.jobs[ji];
  const int tid = threadIdx.x;
  const int Rx = a.p.range_x, Ry = a.p.range_y;
  const int NX = 2 * Rx + 1, NY = 2 * Ry + 1, NC = NX * NY;
  const int cx = job.cmv.col / 8, cy = job.cmv.row / 8;  // truncating, like `cmv.col / 8` in the reference
  const int winW = 64 + 2 * Rx, winH = 64 + 2 * Ry;

  uint32_t *s_org = (uint32_t *)smem;
  uint8_t *s_ref = smem + 64 * 64 * PB;
  uint32_t *s_x = (uint32_t *)(s_ref + (size_t)winH * a.rstride);

  // ---- stage the org region and the reference window (coordinates clamped into
  //      the allocation: pixels outside it are never part of a legal candidate) ----
  {
    const int oxmin = -a.org.xorigin, oxmax = a.org.alloc_w - a.org.xorigin - 1;
    const int oymin = -a.org.yorigin, oymax = a.org.alloc_h - a.org.yorigin - 1;
    for (int i = tid; i < 64 * 64; i += SB_THREADS) {
      const int y = min(max(job.po_y + (i >> 6), oymin), oymax), x = min(max(job.po_x + (i & 63), oxmin), oxmax);
      if (PB == 1) ((uint8_t *)s_org)[i] = ((const uint8_t *)a.org.p0)[(long long)y * a.org.stride + x];
      else ((uint16_t *)s_org)[i] = ((const uint16_t *)a.org.p0)[(long long)y * a.org.stride + x];
    }
    const int rxmin = -a.ref.xorigin, rxmax = a.ref.alloc_w - a.ref.xorigin - 1;
    const int rymin = -a.ref.yorigin, rymax = a.ref.alloc_h - a.ref.yorigin - 1;
    const int x0 = job.po_x + cx - Rx, y0 = job.po_y + cy - Ry;
    const int rowpx = a.rstride / PB;  // includes the slack word
    for (int i = tid; i < winH * rowpx; i += SB_THREADS) {
      const int wy = i / rowpx, wx = i - wy * rowpx;
      const int y = min(max(y0 + wy, rymin), rymax), x = min(max(x0 + wx, rxmin), rxmax);
      if (PB == 1) s_ref[(size_t)wy * a.rstride + wx] = ((const uint8_t *)a.ref.p0)[(long long)y * a.ref.stride + x];
      else ((uint16_t *)(s_ref + (size_t)wy * a.rstride))[wx] = ((const uint16_t *)a.ref.p0)[(long long)y * a.ref.stride + x];
    }
    (void)winW;
  }
  __syncthreads();

  const int c = tid & 15, q = tid >> 4;
  Geo g;
  g.QX = job.po_x + (q & 3) * 16;
  g.QY = job.po_y + (q >> 2) * 16;
  g.W4 = a.p.w_in_b * 4;
  g.H4 = a.p.h_in_b * 4;
  g.cxlo = cx - Rx;
  g.cylo = cy - Ry;
  g.step = a.p.step;
  // every candidate of the nominal window is legal for every block of the region
  // (the most restrictive block is a 4x4 at either edge of the region)
  const bool fast = a.p.step == 1 && (job.po_x + 60 + g.cxlo + 4 >= -16) && (job.po_x + 60 + cx + Rx <= g.W4 + 16) &&
                    (job.po_x + g.cxlo + 4 >= -16) && (job.po_x + cx + Rx <= g.W4 + 16) &&
                    (job.po_y + g.cylo + 4 >= -16) && (job.po_y + 60 + cy + Ry <= g.H4 + 16) &&
                    (job.po_y + 60 + g.cylo + 4 >= -16) && (job.po_y + cy + Ry <= g.H4 + 16);

  unsigned long long best[49], bbest[5];
#pragma unroll
  for (int i = 0; i < 49; i++) best[i] = KEY_MAX;
#pragma unroll
  for (int i = 0; i < 5; i++) bbest[i] = KEY_MAX;

  if (fast) sb_search<PB, true>(a, job, s_org, s_ref, s_x, best, bbest, g, NX, NC, q, c);
  else sb_search<PB, false>(a, job, s_org, s_ref, s_x, best, bbest, g, NX, NC, q, c);

  // ---- reduce over the 16 candidate lanes and write out -----------------------------
  const int qx16 = (q & 3) * 16, qy16 = (q >> 2) * 16;
#pragma unroll
  for (int bi = 0; bi < 49; bi++) {
    unsigned long long key = best[bi];
    int lane = c;
    reduce16(key, lane);
    if (c == 0) {
      int bsize, lx, ly;
      if (bi < 16) { bsize = R1C_BLOCK_4X4; lx = 4 * (bi & 3); ly = 4 * (bi >> 2); }
      else if (bi < 24) { bsize = R1C_BLOCK_4X8; lx = 4 * ((bi - 16) & 3); ly = 8 * ((bi - 16) >> 2); }
      else if (bi < 32) { bsize = R1C_BLOCK_8X4; lx = 8 * ((bi - 24) & 1); ly = 4 * ((bi - 24) >> 1); }
      else if (bi < 36) { bsize = R1C_BLOCK_8X8; lx = 8 * ((bi - 32) & 1); ly = 8 * ((bi - 32) >> 1); }
      else if (bi < 38) { bsize = R1C_BLOCK_8X16; lx = 8 * (bi - 36); ly = 0; }
      else if (bi < 40) { bsize = R1C_BLOCK_16X8; lx = 0; ly = 8 * (bi - 38); }
      else if (bi == 40) { bsize = R1C_BLOCK_16X16; lx = 0; ly = 0; }
      else if (bi < 45) { bsize = R1C_BLOCK_4X16; lx = 4 * (bi - 41); ly = 0; }
      else { bsize = R1C_BLOCK_16X4; lx = 0; ly = 4 * (bi - 45); }
      write_result(a, job, ji, bsize, qx16 + lx, qy16 + ly, key, lane, NX, g);
    }
  }
#pragma unroll
  for (int rd = 0; rd < 5; rd++) {
    const int p = tid + rd * SB_THREADS;
    unsigned long long key = p < 65 * 16 ? bbest[rd] : KEY_MAX;
    int lane = c;
    reduce16(key, lane);
    if (c == 0 && p < 65 * 16) {
      const BigBlk bb = c_big[p >> 4];
      write_result(a, job, ji, bb.bsize, bb.lx * 4, bb.ly * 4, key, lane, NX, g);
    }
  }
}

// host mirror of the c_big table
void build_big_table(BigBlk *t) {
  int n = 0;
  auto add = [&](int bsize, int part_mode) {
    const int bw = blk_w_of(bsize), bh = blk_h_of(bsize);
    for (int y = 0; y < 64; y += bh)
      for (int x = 0; x < 64; x += bw) {
        BigBlk b;
        b.bsize = (unsigned char)bsize;
        b.qx0 = (unsigned char)(x / 16);
        b.qy0 = (unsigned char)(y / 16);
        b.nqx = (unsigned char)(bw >= 16 ? bw / 16 : 1);
        b.nqy = (unsigned char)(bh >= 16 ? bh / 16 : 1);
        b.part = 0;
        if (part_mode == 1) b.part = (unsigned char)(1 + ((x / 8) & 1));  // 8 wide: 8x16 left / right halves
        if (part_mode == 2) b.part = (unsigned char)(3 + ((y / 8) & 1));  // 8 high: 16x8 top / bottom halves
        b.lx = (unsigned char)(x / 4);
        b.ly = (unsigned char)(y / 4);
        t[n++] = b;
      }
  };
  add(R1C_BLOCK_16X32, 0);
  add(R1C_BLOCK_32X16, 0);
  add(R1C_BLOCK_32X32, 0);
  add(R1C_BLOCK_32X64, 0);
  add(R1C_BLOCK_64X32, 0);
  add(R1C_BLOCK_64X64, 0);
  add(R1C_BLOCK_8X32, 1);
  add(R1C_BLOCK_32X8, 2);
  add(R1C_BLOCK_16X64, 0);
  add(R1C_BLOCK_64X16, 0);
}

int launch_sb(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_sb_params *p, int bit_depth,
              const r1c_sb_job *d_jobs, int64_t n, r1c_me_result *d_results) {
  if (n == 0) return R1C_OK;
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  {
    int handled = 0;
    const int rc2 = r1c_sb2_search(ctx, org, ref, p, bit_depth, d_jobs, n, d_results, nullptr, 0, &handled);
    if (rc2 || handled) return rc2;
  }
  static bool table_done[64] = {false};
  if (!table_done[ctx->device & 63]) {
    BigBlk t[65];
    build_big_table(t);
    R1C_CUDA(ctx, cudaMemcpyToSymbol(c_big, t, sizeof t));
    table_done[ctx->device & 63] = true;
  }
  const int PB = org->px_bytes;
  SbArgs a;
  a.org = make_view(org);
  a.ref = make_view(ref);
  a.p = *p;
  a.jobs = d_jobs;
  a.results = d_results;
  a.n = n;
  const int winW = 64 + 2 * p->range_x, winH = 64 + 2 * p->range_y;
  a.rstride = ((winW * PB + 3) & ~3) + 4;
  const size_t smem = (size_t)64 * 64 * PB + (size_t)winH * a.rstride + 2 * 16 * 5 * 16 * 4;
  R1C_REQUIRE(ctx, smem <= 220 * 1024, R1C_ERR_ARG, "search range %dx%d needs %zu bytes of shared memory", p->range_x, p->range_y, smem);
  r1c_prof_begin(ctx, R1C_PROF_ME_SB);
  if (PB == 1) {
    R1C_CUDA(ctx, cudaFuncSetAttribute(me_sb_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    me_sb_kernel<1><<<(unsigned)n, SB_THREADS, smem, ctx->stream>>>(a);
  } else {
    R1C_CUDA(ctx, cudaFuncSetAttribute(me_sb_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    me_sb_kernel<2><<<(unsigned)n, SB_THREADS, smem, ctx->stream>>>(a);
  }
  r1c_prof_end(ctx, R1C_PROF_ME_SB);
  R1C_LAUNCH_CHECK(ctx);
  return R1C_OK;
}

int check_sb(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref, const r1c_sb_params *p, int bit_depth,
             int64_t n) {
  if (!ctx || !org || !ref || !p) return R1C_ERR_ARG;
  R1C_REQUIRE(ctx, org->px_bytes == ref->px_bytes, R1C_ERR_ARG, "org/ref pixel types differ");
  R1C_REQUIRE(ctx, bit_depth_ok(bit_depth, org->px_bytes), R1C_ERR_ARG, "bit_depth %d vs %d-byte pixels", bit_depth, org->px_bytes);
  R1C_REQUIRE(ctx, p->range_x >= 0 && p->range_y >= 0 && (p->step == 1 || p->step == 2), R1C_ERR_ARG, "bad range/step");
  R1C_REQUIRE(ctx, p->w_in_b > 0 && p->h_in_b > 0, R1C_ERR_ARG, "bad frame size");
  R1C_REQUIRE(ctx, ((long long)(2 * p->range_x + 1) * (2 * p->range_y + 1) + 15) / 16 < (1 << IT_BITS), R1C_ERR_ARG, "window too large");
  R1C_REQUIRE(ctx, n >= 0 && n < (1ll << 31), R1C_ERR_ARG, "bad job count");
  return R1C_OK;
}

}  // namespace

extern "C" int r1c_sb_result_offset(int bsize) {
  if (bsize < 0 || bsize >= R1C_BLOCK_SIZES_ALL || sb_count(bsize) == 0) return -1;
  return sb_offset(bsize);
}
extern "C" int r1c_sb_result_count(int bsize) {
  if (bsize < 0 || bsize >= R1C_BLOCK_SIZES_ALL) return 0;
  return sb_count(bsize);
}

extern "C" int r1c_me_sb_batch_dev(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                                   const r1c_sb_params *params, const r1c_sb_job *d_jobs, int64_t n,
                                   int bit_depth, r1c_me_result *d_results) {
  int rc = check_sb(ctx, org, ref, params, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (d_jobs && d_results), R1C_ERR_ARG, "null job/result array");
  return launch_sb(ctx, org, ref, params, bit_depth, d_jobs, n, d_results);
}

extern "C" int r1c_me_sb_batch(r1c_ctx *ctx, const r1c_plane *org, const r1c_plane *ref,
                               const r1c_sb_params *params, const r1c_sb_job *jobs, int64_t n,
                               int bit_depth, r1c_me_result *results) {
  int rc = check_sb(ctx, org, ref, params, bit_depth, n);
  if (rc) return rc;
  R1C_REQUIRE(ctx, n == 0 || (jobs && results), R1C_ERR_ARG, "null job/result array");
  if (n == 0) return R1C_OK;
  const int W4 = params->w_in_b * 4, H4 = params->h_in_b * 4;
  for (int64_t i = 0; i < n; i++) {
    const r1c_sb_job &j = jobs[i];
    R1C_REQUIRE(ctx, (j.po_x & 3) == 0 && (j.po_y & 3) == 0 && j.po_x >= 0 && j.po_y >= 0 && j.po_x < W4 && j.po_y < H4,
                R1C_ERR_ARG, "sb job %lld: region origin must be a multiple of 4 inside the frame", (long long)i);
    // the search centre must be a legal MV for every block of the region, so that each
    // block's window is a sub-window of the nominal one (cmv = 0 always qualifies)
    const int cx = j.cmv.col / 8, cy = j.cmv.row / 8;
    const int bx_max = j.po_x + 60 < W4 - 4 ? j.po_x + 60 : W4 - 4;  // last block origin that exists in the frame
    const int by_max = j.po_y + 60 < H4 - 4 ? j.po_y + 60 : H4 - 4;
    R1C_REQUIRE(ctx, cx >= -(j.po_x + 16 + 4) && cx <= W4 - bx_max + 16 && cy >= -(j.po_y + 16 + 4) && cy <= H4 - by_max + 16,
                R1C_ERR_ARG, "sb job %lld: search centre outside the legal MV range of the region", (long long)i);
  }
  void *h_jobs, *h_res, *d_jobs, *d_res;
  const size_t jb = (size_t)n * sizeof(r1c_sb_job), rb = (size_t)n * R1C_SB_RESULTS * sizeof(r1c_me_result);
  R1C_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((rc = r1c_stage_host(ctx, 0, jb, &h_jobs))) return rc;
  if ((rc = r1c_stage_host(ctx, 1, rb, &h_res))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 0, jb, &d_jobs))) return rc;
  if ((rc = r1c_scratch_dev(ctx, 1, rb, &d_res))) return rc;
  memcpy(h_jobs, jobs, jb);
  memcpy(h_res, results, rb);  // entries the kernel does not emit keep the caller's values
  R1C_CUDA(ctx, cudaMemcpyAsync(d_jobs, h_jobs, jb, cudaMemcpyHostToDevice, ctx->stream));
  R1C_CUDA(ctx, cudaMemcpyAsync(d_res, h_res, rb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = launch_sb(ctx, org, ref, params, bit_depth, (const r1c_sb_job *)d_jobs, n, (r1c_me_result *)d_res))) return rc;
  R1C_CUDA(ctx, cudaMemcpyAsync(h_res, d_res, rb, cudaMemcpyDeviceToHost, ctx->stream));
  R1C_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(results, h_res, rb);
  return R1C_OK;
}
