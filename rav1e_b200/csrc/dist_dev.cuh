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

// SAD of a w x h block, one warp; rows at byte strides so / sr
template <int PB>
__device__ __forceinline__ uint32_t sad_warp(const uint8_t *po, int so, const uint8_t *pr, int sr, int w, int h,
                                            int lane) {
  const int wpr = (w * PB) >> 2;   // 32-bit words per row; w*PB is a multiple of 4
  const int lw = 31 - __clz(wpr);  // wpr is a power of two
  const int total = wpr * h;
  uint32_t acc = 0;
#pragma unroll 4
  for (int g = lane; g < total; g += 32) {
    const int r = g >> lw, c = (g & (wpr - 1)) << 2;
    const uint32_t x = ld4_any(po + r * so + c);
    const uint32_t y = ld4_any(pr + r * sr + c);
    if (PB == 1) {
      acc = sad4_acc(x, y, acc);
    } else {
      const int d0 = (int)(x & 0xffff) - (int)(y & 0xffff);
      const int d1 = (int)(x >> 16) - (int)(y >> 16);
      acc += (uint32_t)abs(d0) + (uint32_t)abs(d1);
    }
  }
  return __reduce_add_sync(0xffffffffu, acc);
}

// one row of `S` differences, row `r` of the tile whose top-left is (po, pr)
template <int PB, int S>
__device__ __forceinline__ void load_diff_row(const uint8_t *po, const uint8_t *pr, int (&d)[S]) {
  if (PB == 1) {
#pragma unroll
    for (int k = 0; k < S; k += 4) {
      uint32_t x = ld4_any(po + k), y = ld4_any(pr + k);
#pragma unroll
      for (int b = 0; b < 4; b++) d[k + b] = (int)((x >> (8 * b)) & 0xff) - (int)((y >> (8 * b)) & 0xff);
    }
  } else {
#pragma unroll
    for (int k = 0; k < S; k += 2) {
      uint32_t x = ld4_any(po + 2 * k), y = ld4_any(pr + 2 * k);
      d[k] = (int)(x & 0xffff) - (int)(y & 0xffff);
      d[k + 1] = (int)(x >> 16) - (int)(y >> 16);
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
__device__ __forceinline__ void hadamard_lanes(int (&d)[S], int lane) {
#pragma unroll
  for (int m = 1; m < S; m <<= 1) {
    const bool upper = (lane & m) != 0;
#pragma unroll
    for (int i = 0; i < S; i++) {
      int o = __shfl_xor_sync(0xffffffffu, d[i], m);
      d[i] = upper ? (o - d[i]) : (d[i] + o);
    }
  }
}

// S = 4 or 8 (tile edge).  32/S tiles are in flight per warp iteration.
template <int PB, int S>
__device__ __forceinline__ uint64_t satd_warp(const uint8_t *po, int so, const uint8_t *pr, int sr,
                                              int w, int h, int lane) {
  constexpr int TPW = 32 / S;  // tiles per warp pass
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
    hadamard_lanes<S>(d, lane);
#pragma unroll
    for (int i = 0; i < S; i++) acc += (uint32_t)abs(d[i]);
  }
  // per-lane partial fits u32 (<= 8 * 64*4095 * 64 tiles); the warp total may not
  uint64_t s = acc;
#pragma unroll
  for (int m = 16; m; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
  return s;
}

// get_satd_ref's tile rule + normalisation (src/dist.rs:287, 326-327)
template <int PB>
__device__ __forceinline__ uint32_t satd_block_warp(const uint8_t *po, int so, const uint8_t *pr, int sr, int w,
                                                    int h, int lane) {
  uint64_t sum;
  int ln;
  if (w == 4 || h == 4) {
    sum = satd_warp<PB, 4>(po, so, pr, sr, w, h, lane);
    ln = 2;
  } else {
    sum = satd_warp<PB, 8>(po, so, pr, sr, w, h, lane);
    ln = 3;
  }
  return (uint32_t)((sum + ((1ull << ln) >> 1)) >> ln);
}

}  // namespace distdev
