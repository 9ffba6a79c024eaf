// itx1d.cuh -- 1-D inverse transforms on register arrays (reference src/transform/inverse.rs):
//   idct<N>      av1_idct4 :35-55, av1_idct8 :118-165, av1_idct16 :261-356, av1_idct32 :541-839, av1_idct64 :848-1544
//   iadst4       av1_iadst4 :63-109            iadst<N>   av1_iadst8 :173-252, av1_iadst16 :364-532
//   iident<N>    av1_iidentity4 / 8 / 16 / 32 :111-116, :254-259, :534-539, :841-846
// Every loop below has compile-time bounds and is fully unrolled, so the arrays live in registers, the
// rotation weights become constant-bank operands and the index arithmetic disappears.  The DCT works in
// place on the bit-reversed input, the way the 32- and 64-point listings of the reference do: position 0/1
// is the c32 pair, positions [M, 2M) are the odd half of the 2M-point transform for M = 2, 4, .. N/2, and
// the add / sub combines then run from the smallest transform outwards.
#pragma once

#include "common.cuh"

namespace itx {

enum { I1_DCT = 0, I1_ADST = 1, I1_ID = 2 };

// COSPI_INV, inverse.rs:22-29 = round(4096 cos(k pi / 128))
static __constant__ int c_cospi[64] = {
    4096, 4095, 4091, 4085, 4076, 4065, 4052, 4036, 4017, 3996, 3973, 3948, 3920, 3889, 3857, 3822,
    3784, 3745, 3703, 3659, 3612, 3564, 3513, 3461, 3406, 3349, 3290, 3229, 3166, 3102, 3035, 2967,
    2896, 2824, 2751, 2675, 2598, 2520, 2440, 2359, 2276, 2191, 2106, 2019, 1931, 1842, 1751, 1660,
    1567, 1474, 1380, 1285, 1189, 1092, 995,  897,  799,  700,  601,  501,  401,  301,  201,  101};

// half_btf with bit = INV_COS_BIT = 12, src/transform/mod.rs:478-488 (wrapping)
__device__ __forceinline__ int btf(int w0, int a, int w1, int b) {
  return (int)((unsigned)w0 * (unsigned)a + (unsigned)w1 * (unsigned)b + 2048u) >> 12;
}
// (cos k, sin k) rotations in the sign patterns the listings use; k in 1/128 turns of pi
__device__ __forceinline__ void rot_a(int k, int &u, int &v) {  // [c(64-k) u - c(k) v, c(k) u + c(64-k) v]
  const int a = u, b = v;
  u = btf(c_cospi[64 - k], a, -c_cospi[k], b);
  v = btf(c_cospi[k], a, c_cospi[64 - k], b);
}
__device__ __forceinline__ void rot_p(int k, int &u, int &v) {  // [-c(k) u + c(64-k) v, c(64-k) u + c(k) v]
  const int a = u, b = v;
  u = btf(-c_cospi[k], a, c_cospi[64 - k], b);
  v = btf(c_cospi[64 - k], a, c_cospi[k], b);
}
__device__ __forceinline__ void rot_q(int k, int &u, int &v) {  // [-c(64-k) u - c(k) v, -c(k) u + c(64-k) v]
  const int a = u, b = v;
  u = btf(-c_cospi[64 - k], a, -c_cospi[k], b);
  v = btf(-c_cospi[k], a, c_cospi[64 - k], b);
}
__device__ __forceinline__ void rot_s(int k, int &u, int &v) {  // [c(k) u + c(64-k) v, c(64-k) u - c(k) v]
  const int a = u, b = v;
  u = btf(c_cospi[k], a, c_cospi[64 - k], b);
  v = btf(c_cospi[64 - k], a, -c_cospi[k], b);
}
__device__ __forceinline__ void rot_t(int k, int &u, int &v) {  // [-c(64-k) u + c(k) v, c(k) u + c(64-k) v]
  const int a = u, b = v;
  u = btf(-c_cospi[64 - k], a, c_cospi[k], b);
  v = btf(c_cospi[k], a, c_cospi[64 - k], b);
}

// clamp_value(v, range), src/transform/mod.rs:492-496, with the two bounds precomputed
struct Range {
  int lo, hi;
};
__device__ __forceinline__ int cl(int v, Range r) { return min(max(v, r.lo), r.hi); }
// The 1-D transforms take CL = false when their clamps are provably no-ops.  Every value the reference clamps
// inside av1_idct* / av1_iadst* is a linear combination of the inputs with weights of magnitude <= 1 (measured on the
// reference's own text, unit inputs through every clamp site: 0.924 .. 0.9999) plus the rounding of at most 162
// half_btf calls, each of which reaches a later clamp site with gain <= 1 (measured the same way): so
//   sum |input| + 128 <= range.hi   ==>   no clamp_value of that transform changes its argument.
constexpr int NOCLAMP_MARGIN = 128;
template <bool CL>
__device__ __forceinline__ int clq(int v, Range r) { return CL ? cl(v, r) : v; }

__host__ __device__ constexpr int bitrev(int v, int bits) {
  int r = 0;
  for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
  return r;
}
__host__ __device__ constexpr int log2c(int n) {
  int l = 0;
  while ((1 << l) < n) l++;
  return l;
}

// odd half of the 2M-point DCT, in place on s[M .. 2M)
template <int N, int M, bool CL>
__device__ __forceinline__ void idct_odd(int (&s)[N], Range rg) {
  constexpr int LM = log2c(M);
#pragma unroll
  for (int i = 0; i < M / 2; i++) rot_a((2 * bitrev(i, LM) + 1) * 32 / M, s[M + i], s[2 * M - 1 - i]);
#pragma unroll
  for (int l = 1; l < LM; l++) {
    const int g = 1 << l;
#pragma unroll
    for (int i = 0; i < M; i++) {  // add / sub inside groups of g, odd groups mirrored
      const int j = i ^ (g - 1);   // partner inside the group
      if (j < i) continue;
      const int u = s[M + i], v = s[M + j];
      if ((i >> l) & 1) {
        s[M + i] = clq<CL>(v - u, rg);
        s[M + j] = clq<CL>(u + v, rg);
      } else {
        s[M + i] = clq<CL>(u + v, rg);
        s[M + j] = clq<CL>(u - v, rg);
      }
    }
#pragma unroll
    for (int i = 0; i < M / 2; i++) {  // rotate the middle half of every block of 2g against its mirror
      const int m = i & (2 * g - 1);
      if (m < g / 2 || m >= g + g / 2) continue;
      int k = 32;
      if (4 * g <= M) k = (64 * g / M) * (2 * bitrev(i >> (l + 1), LM - l - 1) + 1);
      if (m < g) rot_p(k, s[M + i], s[2 * M - 1 - i]);
      else rot_q(k, s[M + i], s[2 * M - 1 - i]);
    }
  }
}

template <int N, int M, bool CL>
__device__ __forceinline__ void idct_odd_all(int (&s)[N], Range rg) {
  idct_odd<N, M, CL>(s, rg);
  if constexpr (2 * M < N) idct_odd_all<N, 2 * M, CL>(s, rg);
}

template <int N, bool CL = true>
__device__ __forceinline__ void idct(const int (&in)[N], int (&out)[N], Range rg) {
  constexpr int LN = log2c(N);
  int s[N];
#pragma unroll
  for (int p = 0; p < N; p++) s[p] = in[bitrev(p, LN)];
  {
    const int u = s[0], v = s[1];
    s[0] = btf(c_cospi[32], u, c_cospi[32], v);
    s[1] = btf(c_cospi[32], u, -c_cospi[32], v);
  }
  idct_odd_all<N, 2, CL>(s, rg);
#pragma unroll
  for (int l = 1; l < LN; l++) {
    const int M = 1 << l;
#pragma unroll
    for (int i = 0; i < M; i++) {
      const int u = s[i], v = s[2 * M - 1 - i];
      s[i] = clq<CL>(u + v, rg);
      s[2 * M - 1 - i] = clq<CL>(u - v, rg);
    }
  }
#pragma unroll
  for (int p = 0; p < N; p++) out[p] = s[p];
}

// round_shift(v, 12) of a wrapping i32 value (unsigned arithmetic: overflow is defined and wraps like the
// reference's release build)
__device__ __forceinline__ int rshift12(unsigned v) { return (int)(v + 2048u) >> 12; }

// av1_iadst4: SINPI_INV = {0, 1321, 2482, 3344, 3803}, :31; plain i32 products, one rounding at the end
__device__ __forceinline__ void iadst4(const int (&in)[4], int (&out)[4]) {
  const unsigned x0 = (unsigned)in[0], x1 = (unsigned)in[1], x2 = (unsigned)in[2], x3 = (unsigned)in[3];
  const unsigned p = 1321u * x0 + 3803u * x2 + 2482u * x3;
  const unsigned q = 2482u * x0 - 1321u * x2 - 3803u * x3;
  const unsigned m = 3344u * x1;
  out[0] = rshift12(p + m);
  out[1] = rshift12(q + m);
  out[2] = rshift12(3344u * (x0 - x2 + x3));
  out[3] = rshift12(p + q - m);
}

template <int N, bool CL = true>
__device__ __forceinline__ void iadst(const int (&in)[N], int (&out)[N], Range rg) {
  constexpr int LN = log2c(N);
  int s[N];
#pragma unroll
  for (int i = 0; i < N; i += 2) {  // stages 1 and 2
    s[i] = in[N - 1 - i];
    s[i + 1] = in[i];
    rot_s((2 * i + 1) * 32 / N, s[i], s[i + 1]);
  }
#pragma unroll
  for (int l = LN - 1; l >= 1; l--) {
    const int h = 1 << l;
#pragma unroll
    for (int i = 0; i < N; i++) {
      if (i & h) continue;
      const int u = s[i], v = s[i + h];
      s[i] = clq<CL>(u + v, rg);
      s[i + h] = clq<CL>(u - v, rg);
    }
#pragma unroll
    for (int i = 0; i < N; i += 2) {  // pairs of the upper half of every block of 2h
      if (!(i & h)) continue;
      if (h == 2) rot_s(32, s[i], s[i + 1]);
      else {
        const int k = (4 * ((i & (h / 2 - 1)) >> 1) + 1) * 64 / h;
        if (i & (h / 2)) rot_t(k, s[i], s[i + 1]);
        else rot_s(k, s[i], s[i + 1]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < N; i++) {  // Gray-code order, bit-reversed; odd outputs negated
    const int v = s[bitrev(i ^ (i >> 1), LN)];
    out[i] = (i & 1) ? -v : v;
  }
}

template <int N>
__device__ __forceinline__ void iident(const int (&in)[N], int (&out)[N]) {
#pragma unroll
  for (int i = 0; i < N; i++)
    out[i] = N == 4 ? rshift12(5793u * (unsigned)in[i]) : N == 8 ? (int)(2u * (unsigned)in[i])
           : N == 16 ? rshift12(5793u * 2u * (unsigned)in[i]) : (int)(4u * (unsigned)in[i]);
}

template <int TYPE, int N, bool CL = true>
__device__ __forceinline__ void inv1d(const int (&in)[N], int (&out)[N], Range rg) {
  if constexpr (TYPE == I1_DCT) idct<N, CL>(in, out, rg);
  else if constexpr (TYPE == I1_ADST) {
    if constexpr (N == 4) iadst4(in, out);
    else iadst<N, CL>(in, out, rg);
  } else iident<N>(in, out);
}
// does the transform clamp at all?  (identity and the 4-point ADST do not)
template <int TYPE, int N>
constexpr bool inv1d_clamps() { return TYPE == I1_DCT || (TYPE == I1_ADST && N > 4); }

}  // namespace itx
