// me_common.cuh -- helpers shared by the motion-search kernels (reference src/me.rs).
#pragma once

#include "common.cuh"

namespace me {

// get_mv_rate, src/me.rs:1005-1021 (i16 arithmetic; `diff >> 1` unless high precision)
__host__ __device__ inline uint32_t diff_to_rate(int16_t diff, int allow_hp) {
  const int16_t d = allow_hp ? diff : (int16_t)(diff >> 1);
  if (d == 0) return 0;
  const uint32_t a = (uint32_t)(uint16_t)(d < 0 ? -d : d);
#ifdef __CUDA_ARCH__
  return 2u * (uint32_t)(32 - __clz((int)a));  // 16 - leading_zeros(u16)
#else
  int bits = 0;
  for (uint32_t v = a; v; v >>= 1) bits++;
  return 2u * (uint32_t)bits;
#endif
}
__host__ __device__ inline uint32_t mv_rate(r1c_mv a, r1c_mv b, int allow_hp) {
  return diff_to_rate((int16_t)(a.row - b.row), allow_hp) + diff_to_rate((int16_t)(a.col - b.col), allow_hp);
}
// rate = min(rate(mv, pmv0), rate(mv, pmv1) + 1), src/me.rs:851-853, 979-981
__host__ __device__ inline uint32_t mv_rate2(r1c_mv mv, r1c_mv p0, r1c_mv p1, int allow_hp) {
  const uint32_t r1 = mv_rate(mv, p0, allow_hp), r2 = mv_rate(mv, p1, allow_hp) + 1;
  return r1 < r2 ? r1 : r2;
}

__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t acc) {
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(acc) : "r"(a), "r"(b), "r"(acc));
  return acc;
}
// two u16 lanes: |a - b| per half as max - min (two VIMNMX.U16x2 and a subtraction that cannot borrow), both halves
// added to acc by one IDP.2A with unit weights.  Exact for any u16 values.  (The scalar `vabsdiff.u32 ... .add` of
// round 1 is emulated on sm_100a: 7 instructions per pixel.)
__device__ __forceinline__ uint32_t sad2_acc(uint32_t a, uint32_t b, uint32_t acc) {
  return __dp2a_lo(__vmaxu2(a, b) - __vminu2(a, b), 0x0101u, acc);
}
template <int PB>
__device__ __forceinline__ uint32_t sad_word(uint32_t a, uint32_t b, uint32_t acc) {
  return PB == 1 ? sad4_acc(a, b, acc) : sad2_acc(a, b, acc);
}

// 4 bytes at an arbitrary byte address of GLOBAL memory from aligned words only
__device__ __forceinline__ uint32_t ldg4_any(const uint8_t *p) {
  const uintptr_t a = (uintptr_t)p;
  const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(a & 3) * 8;
  const uint32_t lo = __ldg(w);
  const uint32_t hi = sh ? __ldg(w + 1) : 0u;
  return __funnelshift_r(lo, hi, sh);
}

}  // namespace me
