// dist_dev.cuh -- device building blocks of SAD / SATD shared by dist.cu and
// subpel.cu (reference semantics: get_sad_ref src/dist.rs:25-46, get_satd_ref :197-328).
#pragma once

#include "common.cuh"

namespace distdev {

// 4 bytes at an arbitrary byte address, built from aligned words only (never
// touches a word that holds none of the requested bytes -> cannot leave the
// allocation).
__device__ __forceinline__ uint32_t ld4_any(const uint8_t *p) {
  uintptr_t a = (uintptr_t)p;
  const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
  uint32_t sh = (uint32_t)(a & 3) * 8;
  uint32_t lo = __ldg(w);
  uint32_t hi = sh ? __ldg(w + 1) : 0u;
  return __funnelshift_r(lo, hi, sh);
}

__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t acc) {
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(acc) : "r"(a), "r"(b), "r"(acc));
  return acc;
}

// NW consecutive 32-bit words starting at an arbitrary byte address: NW (+1 when the
// address is not word aligned) aligned loads and a funnel shift each.  Like ld4_any, no
// word is touched that holds none of the requested bytes.
template <int NW>
__device__ __forceinline__ void ld_words(const uint8_t *p, uint32_t (&out)[NW]) {
  const uintptr_t a = (uintptr_t)p;
  const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(a & 3) * 8;
  uint32_t v[NW + 1];
#pragma unroll
  for (int i = 0; i < NW; i++) v[i] = __ldg(w + i);
  v[NW] = sh ? __ldg(w + NW) : 0u;
#pragma unroll
  for (int i = 0; i < NW; i++) out[i] = __funnelshift_r(v[i], v[i + 1], sh);
}

// The same NW words fetched with NW*4-byte vector loads from the enclosing aligned
// chunks (one load when the address is aligned, else two) -- a warp whose lanes read
// neighbouring tiles then consumes whole sectors per instruction instead of touching each
// sector NW times.  Only legal when the plane's allocation starts and ends on a multiple of
// NW*4 bytes (checked by the launcher): the second chunk is read only if it holds requested
// bytes, so no load leaves the allocation.  NW = 2 or 4.
template <int NW>
__device__ __forceinline__ void ld_words_vec(const uint8_t *p, uint32_t (&out)[NW]) {
  constexpr uintptr_t VB = NW * 4;
  const uintptr_t a = (uintptr_t)p;
  const uintptr_t a0 = a & ~(VB - 1);
  const uint32_t off = (uint32_t)(a & (VB - 1));
  uint32_t v[2 * NW + 1];
  if (NW == 4) {
    const uint4 lo = __ldg((const uint4 *)a0);
    uint4 hi = make_uint4(0, 0, 0, 0);
    if (off) hi = __ldg((const uint4 *)(a0 + VB));
    v[0] = lo.x, v[1] = lo.y, v[2] = lo.z, v[3] = lo.w, v[4] = hi.x, v[5] = hi.y, v[6] = hi.z, v[7] = hi.w;
  } else {
    const uint2 lo = __ldg((const uint2 *)a0);
    uint2 hi = make_uint2(0, 0);
    if (off) hi = __ldg((const uint2 *)(a0 + VB));
    v[0] = lo.x, v[1] = lo.y, v[2] = hi.x, v[3] = hi.y;
  }
  v[2 * NW] = 0;
  // drop off/4 whole words with a log-shifter of selects, then the byte remainder
  const uint32_t wo = off >> 2, sh = (off & 3) * 8;
  if (NW == 4) {
    if (wo & 2) {
#pragma unroll
      for (int i = 0; i < 2 * NW - 1; i++) v[i] = v[i + 2];
    }
  }
  if (wo & 1) {
#pragma unroll
    for (int i = 0; i < 2 * NW; i++) v[i] = v[i + 1];
  }
#pragma unroll
  for (int i = 0; i < NW; i++) out[i] = __funnelshift_r(v[i], v[i + 1], sh);
}

template <int NW, bool VEC>
__device__ __forceinline__ void ld_row(const uint8_t *p, uint32_t (&out)[NW]) {
  if (VEC && NW >= 2) ld_words_vec<NW>(p, out);
  else ld_words<NW>(p, out);
}

// two u16 lanes: |a - b| per half as max - min (two VIMNMX.U16x2 and a subtraction that cannot borrow), both halves
// added to acc by one IDP.2A with unit weights.  Exact for any u16 values.  (The scalar `vabsdiff.u32 ... .add` of
// round 1 is emulated on sm_100a: 7 instructions per pixel.)
__device__ __forceinline__ uint32_t sad2_acc(uint32_t a, uint32_t b, uint32_t acc) {
  return __dp2a_lo(__vmaxu2(a, b) - __vminu2(a, b), 0x0101u, acc);
}

// rows of 4 * NW bytes per lane and step
template <int PB, int G, int NW>
__device__ __forceinline__ uint32_t sad_rows(const uint8_t *po, int so, const uint8_t *pr, int sr, int w, int h,
                                            int lane) {
  const int cpr = (w * PB) / (4 * NW);  // chunks per row, a power of two
  const int lc = 31 - __clz(cpr);
  const int total = cpr * h;
  uint32_t acc = 0;
#pragma unroll 2
  for (int g = lane; g < total; g += G) {
    const int r = g >> lc, c = (g & (cpr - 1)) * (4 * NW);
    uint32_t x[NW], y[NW];
    ld_words<NW>(po + r * so + c, x);
    ld_words<NW>(pr + r * sr + c, y);
#pragma unroll
    for (int i = 0; i < NW; i++) {
      if (PB == 1) {
        acc = sad4_acc(x[i], y[i], acc);
      } else {
        acc = sad2_acc(x[i], y[i], acc);
      }
    }
  }
  return acc;
}

// SAD of a w x h block by a group of G consecutive lanes (G = 32: one warp; smaller
// groups let a warp work on 32 / G jobs at once, `mask` = the group's lanes, `lane` =
// index inside the group); rows at byte strides so / sr
template <int PB, int G = 32>
__device__ __forceinline__ uint32_t sad_warp(const uint8_t *po, int so, const uint8_t *pr, int sr, int w, int h,
                                            int lane, unsigned mask = 0xffffffffu) {
  const int bw = w * PB;  // bytes per row: 4 .. 256
  uint32_t acc;
  if (bw >= 16) acc = sad_rows<PB, G, 4>(po, so, pr, sr, w, h, lane);
  else if (bw == 8) acc = sad_rows<PB, G, 2>(po, so, pr, sr, w, h, lane);
  else acc = sad_rows<PB, G, 1>(po, so, pr, sr, w, h, lane);
  if (G == 32) return __reduce_add_sync(mask, acc);
#pragma unroll
  for (int m = G >> 1; m; m >>= 1) acc += __shfl_xor_sync(mask, acc, m);
  return acc;
}

// one row of `S` differences, row `r` of the tile whose top-left is (po, pr)
template <int PB, int S, bool VEC = false>
__device__ __forceinline__ void load_diff_row(const uint8_t *po, const uint8_t *pr, int (&d)[S]) {
  constexpr int NW = S * PB / 4;
  uint32_t x[NW], y[NW];
  ld_row<NW, VEC>(po, x);
  ld_row<NW, VEC>(pr, y);
  if (PB == 1) {
#pragma unroll
    for (int k = 0; k < NW; k++)
#pragma unroll
      for (int b = 0; b < 4; b++) d[4 * k + b] = (int)((x[k] >> (8 * b)) & 0xff) - (int)((y[k] >> (8 * b)) & 0xff);
  } else {
#pragma unroll
    for (int k = 0; k < NW; k++) {
      d[2 * k] = (int)(x[k] & 0xffff) - (int)(y[k] & 0xffff);
      d[2 * k + 1] = (int)(x[k] >> 16) - (int)(y[k] >> 16);
    }
  }
}

template <int S>
__device__ __forceinline__ void hadamard_regs(int (&d)[S]) {
#pragma unroll
  for (int m = 1; m < S; m <<= 1) {
#pragma unroll
    for (int i = 0; i < S; i++) {
      if ((i & m) == 0) {
        int a = d[i], b = d[i | m];
        d[i] = a + b;
        d[i | m] = a - b;
      }
    }
  }
}

template <int S>
__device__ __forceinline__ void hadamard_lanes(int (&d)[S], int lane, unsigned mask = 0xffffffffu) {
#pragma unroll
  for (int m = 1; m < S; m <<= 1) {
    const int sg = (lane & m) ? -1 : 1;  // upper lane of the pair: o - d, lower: d + o -- one IMAD either way
#pragma unroll
    for (int i = 0; i < S; i++) {
      int o = __shfl_xor_sync(mask, d[i], m);
      d[i] = (int)((unsigned)d[i] * (unsigned)sg + (unsigned)o);
    }
  }
}

// S = 4 or 8 (tile edge).  G/S tiles are in flight per pass of a group of G lanes.
template <int PB, int S, int G = 32>
__device__ __forceinline__ uint64_t satd_warp(const uint8_t *po, int so, const uint8_t *pr, int sr,
                                              int w, int h, int lane, unsigned mask = 0xffffffffu) {
  constexpr int TPW = G / S;  // tiles per pass
  const int row = lane & (S - 1), slot = lane / S;
  const int tx = w / S, ntiles = tx * (h / S);
  uint32_t acc = 0;
  for (int t0 = 0; t0 < ntiles; t0 += TPW) {
    int t = t0 + slot;
    int d[S];
    if (t < ntiles) {
      int ty = t / tx, txi = t - ty * tx;
      long long oo = (long long)(ty * S + row) * so + txi * S * PB;
      long long ro = (long long)(ty * S + row) * sr + txi * S * PB;
      load_diff_row<PB, S>(po + oo, pr + ro, d);
    } else {
#pragma unroll
      for (int i = 0; i < S; i++) d[i] = 0;
    }
    hadamard_regs<S>(d);
    hadamard_lanes<S>(d, lane, mask);
#pragma unroll
    for (int i = 0; i < S; i++) acc += (uint32_t)abs(d[i]);
  }
  // per-lane partial fits u32 (<= 8 * 64*4095 * 256 tiles / TPW); the group total may not
  uint64_t s = acc;
#pragma unroll
  for (int m = G >> 1; m; m >>= 1) s += __shfl_xor_sync(mask, s, m);
  return s;
}

// get_satd_ref's tile rule + normalisation (src/dist.rs:287, 326-327)
template <int PB, int G = 32>
__device__ __forceinline__ uint32_t satd_block_warp(const uint8_t *po, int so, const uint8_t *pr, int sr, int w,
                                                    int h, int lane, unsigned mask = 0xffffffffu) {
  uint64_t sum;
  int ln;
  if (w == 4 || h == 4) {
    sum = satd_warp<PB, 4, G>(po, so, pr, sr, w, h, lane, mask);
    ln = 2;
  } else {
    sum = satd_warp<PB, 8, G>(po, so, pr, sr, w, h, lane, mask);
    ln = 3;
  }
  return (uint32_t)((sum + ((1ull << ln) >> 1)) >> ln);
}

}  // namespace distdev
