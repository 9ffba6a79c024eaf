// tx1d.cuh -- 1-D forward transforms as register-resident device code.
//
// What is computed: the integer lifting transforms of the reference's
// src/transform/forward.rs (Daala-derived DCT-II 4..64, DST-IV 8/16, DST-VII 4,
// identity) with exactly its rounding: tx_mul = (x*c + (1<<s>>1)) >> s in
// wrapping i32 (forward.rs:114-117), rshift1 rounds toward zero (:119-121),
// add_avg / sub_avg floor (:123-129).  Constants come from the code, not from
// its comments (several differ).
//
// How it is organised here: every transform works in place on a fully unrolled
// `int v[N]` that lives in registers; an "asymmetric pair" of the reference
// (half, full) is two ints.  All loops have compile-time trip counts.
#pragma once

namespace tx {

struct HF { int h, f; };  // the reference's (half, full) tuple

__device__ __forceinline__ int mulc(int x, int c, int s) {
  // wrapping 32-bit multiply-add, arithmetic shift
  return (int)((unsigned)x * (unsigned)c + (unsigned)((1 << s) >> 1)) >> s;
}
__device__ __forceinline__ int half0(int x) { return (int)((unsigned)x + ((unsigned)x >> 31)) >> 1; }
__device__ __forceinline__ int avg_p(int a, int b) { return (int)((unsigned)a + (unsigned)b) >> 1; }
__device__ __forceinline__ int avg_m(int a, int b) { return (int)((unsigned)a - (unsigned)b) >> 1; }
__device__ __forceinline__ int wadd(int a, int b) { return (int)((unsigned)a + (unsigned)b); }
__device__ __forceinline__ int wsub(int a, int b) { return (int)((unsigned)a - (unsigned)b); }
__device__ __forceinline__ HF hf(int x) { return HF{half0(x), x}; }

// RotateKernel family (forward.rs:201-277).  ADDK: how t = f(p1, p0.h) is formed;
// the output-1 sign and the optional rshift1 of `c` follow from the variant.
enum RotK { R_ADD, R_ADDAVG, R_ADDSHIFT, R_SUB, R_SUBAVG, R_SUBSHIFT };

template <RotK K>
__device__ __forceinline__ void rot_h(HF p0, int p1, int c0, int s0, int c1, int s1, int c2, int s2,
                                      int &o0, int &o1) {
  int t;
  if (K == R_ADD || K == R_ADDSHIFT) t = wadd(p1, p0.h);
  else if (K == R_ADDAVG) t = avg_p(p1, p0.h);
  else if (K == R_SUB || K == R_SUBSHIFT) t = wsub(p1, p0.h);
  else t = avg_m(p1, p0.h);
  const int a = mulc(p0.f, c0, s0), b = mulc(p1, c1, s1), c = mulc(t, c2, s2);
  const int sh = (K == R_ADDSHIFT || K == R_SUBSHIFT) ? half0(c) : c;
  o0 = wadd(b, c);
  o1 = (K == R_ADD || K == R_ADDAVG || K == R_ADDSHIFT) ? wsub(a, sh) : wadd(a, sh);
}
template <RotK K>
__device__ __forceinline__ void rot_f(int p0, int p1, int c0, int s0, int c1, int s1, int c2, int s2,
                                      int &o0, int &o1) {
  rot_h<K>(HF{p0, p0}, p1, c0, s0, c1, s1, c2, s2, o0, o1);
}
// RotateKernelNeg (forward.rs:222-232, 279-285)
template <bool AVG>
__device__ __forceinline__ void rot_n(int p0, int p1, int c0, int s0, int c1, int s1, int c2, int s2,
                                      int &o0, int &o1) {
  const int t = AVG ? avg_m(p0, p1) : wsub(p0, p1);
  const int a = mulc(p0, c0, s0), b = mulc(p1, c1, s1), c = mulc(t, c2, s2);
  o0 = wsub(b, c);
  o1 = wsub(c, a);
}
// RotateKernelPi4 (forward.rs:164-199)
enum Pi4K { Q_ADD, Q_ADDAVG, Q_SUB, Q_SUBAVG };
template <Pi4K K>
__device__ __forceinline__ void rot_q(int p0, int p1, int c0, int s0, int c1, int s1, int &o0, int &o1) {
  int t;
  if (K == Q_ADD) t = wadd(p1, p0);
  else if (K == Q_ADDAVG) t = avg_p(p1, p0);
  else if (K == Q_SUB) t = wsub(p1, p0);
  else t = avg_m(p1, p0);
  const int a = mulc(p0, c0, s0);
  const int r = mulc(t, c1, s1);
  o0 = r;
  o1 = (K == Q_ADD || K == Q_ADDAVG) ? wsub(a, r) : wadd(a, r);
}

// butterflies (forward.rs:287-324)
__device__ __forceinline__ void b_add(int p0, int p1, HF &q0, int &q1h) {
  const int s = wadd(p0, p1), h = half0(s);
  q0 = HF{h, s};
  q1h = wsub(p1, h);
}
__device__ __forceinline__ void b_sub(int p0, int p1, HF &q0, int &q1h) {
  const int s = wsub(p0, p1), h = half0(s);
  q0 = HF{h, s};
  q1h = wadd(p1, h);
}
__device__ __forceinline__ void b_neg(int p0, int p1, int &q0h, HF &q1) {
  const int d = wsub(p0, p1), h = half0(d);
  q0h = wsub(p0, h);
  q1 = HF{h, d};
}
__device__ __forceinline__ void b_add_a(HF p0, int p1h, int &q0, int &q1) {
  const int t = wadd(p1h, p0.h);
  q0 = wsub(p0.f, t);
  q1 = t;
}
__device__ __forceinline__ void b_sub_a(HF p0, int p1h, int &q0, int &q1) {
  const int t = wsub(p1h, p0.h);
  q0 = wadd(p0.f, t);
  q1 = t;
}
__device__ __forceinline__ void b_neg_a(int p0h, HF p1, int &q0, int &q1) {
  const int t = wadd(p0h, p1.h);
  q0 = t;
  q1 = wsub(t, p1.f);
}

// ---------------------------------------------------------------- 2-point ----
__device__ __forceinline__ void dct2_asym(int p0h, HF p1, int &o0, int &o1) { b_neg_a(p0h, p1, o0, o1); }
__device__ __forceinline__ void dst2_asym(HF p0, int p1h, int &o0, int &o1) {
  rot_h<R_ADD>(p0, p1h, 473, 9, 3135, 12, 4433, 13, o0, o1);
}
__device__ __forceinline__ void dct2(int p0, int p1, int &o0, int &o1) {
  // (p1, p0) = Pi4SubAvg(p1, p0); returns (p0, p1)   forward.rs:407-412
  int a, b;
  rot_q<Q_SUBAVG>(p1, p0, 11585, 13, 11585, 13, a, b);
  o0 = b;
  o1 = a;
}
__device__ __forceinline__ void dst2(int p0, int p1, int &o0, int &o1) {
  rot_f<R_ADDAVG>(p0, p1, 10703, 13, 8867, 14, 3135, 12, o0, o1);
}

// ---------------------------------------------------------------- 4-point ----
// forward.rs:349-361; o[] in the kernel's own (un-permuted) order
__device__ __forceinline__ void dct4_core(int q0, int q1, int q2, int q3, int *o) {
  int q0h, q2h;
  HF Q3, Q1;
  b_neg(q0, q3, q0h, Q3);
  b_add(q1, q2, Q1, q2h);
  dct2_asym(q0h, Q1, o[0], o[1]);
  dst2_asym(Q3, q2h, o[3], o[2]);
}
// forward.rs:421-433
__device__ __forceinline__ void dct4_asym(int q0h, HF Q1, int q2h, HF Q3, int *o) {
  int q0, q1, q2, q3;
  b_neg_a(q0h, Q3, q0, q3);
  b_sub_a(Q1, q2h, q1, q2);
  dct2(q0, q1, o[0], o[1]);
  dst2(q3, q2, o[3], o[2]);
}
// stages 1-2 shared by daala_fdst_iv_4 (forward.rs:605-614) and _asym (:456-465)
__device__ __forceinline__ void dst4_tail(int q0, int q1, int q2, int q3, int *o) {
  b_sub_a(hf(q2), q3, q2, q3);
  b_sub_a(hf(q0), q1, q0, q1);
  rot_q<Q_ADDAVG>(q2, q1, 11585, 13, 11585, 13, q2, q1);
  o[0] = q0; o[1] = q1; o[2] = q2; o[3] = q3;
}
// forward.rs:435-466
__device__ __forceinline__ void dst4_asym(HF Q0, int q1h, HF Q2, int q3h, int *o) {
  int q0, q1, q2, q3;
  rot_h<R_ADDSHIFT>(Q0, q3h, 9633, 14, 12873, 13, 12785, 15, q0, q3);
  rot_h<R_SUBSHIFT>(Q2, q1h, 11363, 14, 18081, 15, 4551, 12, q2, q1);
  dst4_tail(q0, q1, q2, q3, o);
}
// forward.rs:590-615
__device__ __forceinline__ void dst4_core(int q0, int q1, int q2, int q3, int *o) {
  rot_f<R_ADDSHIFT>(q0, q3, 13623, 14, 4551, 12, 565, 11, q0, q3);
  rot_f<R_SUBSHIFT>(q2, q1, 16069, 14, 12785, 15, 1609, 11, q2, q1);
  dst4_tail(q0, q1, q2, q3, o);
}

template <int N>
__device__ __forceinline__ void reverse_in_place(int *a) {
#pragma unroll
  for (int i = 0; i < N / 2; i++) {
    int t = a[i];
    a[i] = a[N - 1 - i];
    a[N - 1 - i] = t;
  }
}

// ---------------------------------------------------------------- 8-point ----
// forward.rs:468-481
__device__ __forceinline__ void dct8_core(const int *r, int *o) {
  int r0h, r6h, r2h, r4h;
  HF R7, R1, R5, R3;
  b_neg(r[0], r[7], r0h, R7);
  b_add(r[1], r[6], R1, r6h);
  b_neg(r[2], r[5], r2h, R5);
  b_add(r[3], r[4], R3, r4h);
  dct4_asym(r0h, R1, r2h, R3, o);
  dst4_asym(R7, r6h, R5, r4h, o + 4);
  reverse_in_place<4>(o + 4);
}
// stages 1-3 of the 8-point DST-IV.  ASYM selects the stage-3 constants of
// daala_fdst_iv_8_asym (forward.rs:675-684) vs daala_fdst_iv_8 (:550-559).
template <bool ASYM>
__device__ __forceinline__ void dst8_tail(int r0, int r1, int r2, int r3, int r4, int r5, int r6, int r7, int *o) {
  HF R0, R2, R5, R7;
  int r3h, r1h, r6h, r4h;
  b_add(r0, r3, R0, r3h);
  b_sub(r2, r1, R2, r1h);
  b_add(r5, r6, R5, r6h);
  b_sub(r7, r4, R7, r4h);
  b_add_a(R7, r6h, r7, r6);
  b_add_a(R5, r3h, r5, r3);
  b_add_a(R2, r4h, r2, r4);
  b_sub_a(R0, r1h, r0, r1);
  if (ASYM) {
    rot_f<R_SUBAVG>(r3, r4, 669, 9, 8867, 14, 3135, 12, r3, r4);
    rot_n<true>(r2, r5, 669, 9, 8867, 14, 3135, 12, r2, r5);
    rot_q<Q_SUBAVG>(r1, r6, 5793, 12, 11585, 13, r1, r6);
  } else {
    rot_f<R_SUBAVG>(r3, r4, 10703, 13, 8867, 14, 3135, 12, r3, r4);
    rot_n<true>(r2, r5, 10703, 13, 8867, 14, 3135, 12, r2, r5);
    rot_q<Q_SUBAVG>(r1, r6, 11585, 13, 11585, 13, r1, r6);
  }
  o[0] = r0; o[1] = r1; o[2] = r2; o[3] = r3; o[4] = r4; o[5] = r5; o[6] = r6; o[7] = r7;
}
// forward.rs:509-562
__device__ __forceinline__ void dst8_core(const int *r, int *o) {
  int r0, r1, r2, r3, r4, r5, r6, r7;
  rot_f<R_ADD>(r[0], r[7], 17911, 14, 14699, 14, 803, 13, r0, r7);
  rot_f<R_SUB>(r[6], r[1], 20435, 14, 21845, 15, 1189, 12, r6, r1);
  rot_f<R_ADD>(r[2], r[5], 22173, 14, 3363, 13, 15447, 15, r2, r5);
  rot_f<R_SUB>(r[4], r[3], 23059, 14, 2271, 14, 5197, 13, r4, r3);
  dst8_tail<false>(r0, r1, r2, r3, r4, r5, r6, r7, o);
}
// forward.rs:617-631 : inputs (r0h, r1, r2h, r3, r4h, r5, r6h, r7): h[] at even, P[] at odd
__device__ __forceinline__ void dct8_asym(const int *h, const HF *P, int *o) {
  int r[8];
  b_neg_a(h[0], P[7], r[0], r[7]);
  b_sub_a(P[1], h[6], r[1], r[6]);
  b_neg_a(h[2], P[5], r[2], r[5]);
  b_sub_a(P[3], h[4], r[3], r[4]);
  dct4_core(r[0], r[1], r[2], r[3], o);
  dst4_core(r[7], r[6], r[5], r[4], o + 4);
  reverse_in_place<4>(o + 4);
}
// forward.rs:633-687 : inputs (r0, r1h, r2, r3h, r4, r5h, r6, r7h): P[] at even, h[] at odd
__device__ __forceinline__ void dst8_asym(const HF *P, const int *h, int *o) {
  int r0, r1, r2, r3, r4, r5, r6, r7;
  rot_h<R_ADD>(P[0], h[7], 12665, 14, 5197, 12, 2271, 14, r0, r7);
  rot_h<R_SUB>(P[6], h[1], 14449, 14, 30893, 15, 3363, 13, r6, r1);
  rot_h<R_ADD>(P[2], h[5], 15679, 14, 1189, 11, 5461, 13, r2, r5);
  rot_h<R_SUB>(P[4], h[3], 16305, 14, 803, 12, 14699, 14, r4, r3);
  dst8_tail<true>(r0, r1, r2, r3, r4, r5, r6, r7, o);
}

// The first stage of every even-size DCT (forward.rs:472-475, 694-701,
// 1079-1094, 1559-1567): slot i pairs with N-1-i; even i -> butterfly_neg,
// odd i -> butterfly_add.  Produces for the low half (i < N/2):
//   even i: h[i] = x_i half, P[i] = pair of x_(N-1-i)
//   odd  i: P[i] = pair of x_i, h[i] = x_(N-1-i) half
// which is exactly the argument order the embedded asymmetric DCT (h even, P
// odd) and asymmetric DST (P even, h odd) take.
template <int N>
__device__ __forceinline__ void dct_split(const int *x, int *h, HF *P) {
#pragma unroll
  for (int i = 0; i < N / 2; i += 2) {
    b_neg(x[i], x[N - 1 - i], h[i], P[i]);
    b_add(x[i + 1], x[N - 2 - i], P[i + 1], h[i + 1]);
  }
}
// First stage of the asymmetric DCTs (forward.rs:622-625, 923-930, 1220-1235):
// inputs h[] at even slots / P[] at odd slots over N slots.
template <int N>
__device__ __forceinline__ void dct_asym_merge(const int *h, const HF *P, int *x) {
#pragma unroll
  for (int i = 0; i < N / 2; i += 2) {
    b_neg_a(h[i], P[N - 1 - i], x[i], x[N - 1 - i]);
    b_sub_a(P[i + 1], h[N - 2 - i], x[i + 1], x[N - 2 - i]);
  }
}

// --------------------------------------------------------------- 16-point ----
// forward.rs:689-707
__device__ __forceinline__ void dct16_core(const int *s, int *o) {
  int h[8];
  HF P[8];
  dct_split<16>(s, h, P);
  dct8_asym(h, P, o);
  dst8_asym(P, h, o + 8);
  reverse_in_place<8>(o + 8);
}
// forward.rs:917-936
__device__ __forceinline__ void dct16_asym(const int *h, const HF *P, int *o) {
  int s[16];
  dct_asym_merge<16>(h, P, s);
  dct8_core(s, o);
  int r[8];
#pragma unroll
  for (int i = 0; i < 8; i++) r[i] = s[15 - i];  // (sf, se, sd, sc, sb, sa, s9, s8)
  dst8_core(r, o + 8);
  reverse_in_place<8>(o + 8);
}

// Stages 1-5 of the 16-point DST-IV: daala_fdst_iv_16 (forward.rs:797-872,
// ASYM = false) and daala_fdst_iv_16_asym (:994-1069, ASYM = true) differ only
// in the rotation kernels / constants of stages 3 and 5.
template <bool ASYM>
__device__ __forceinline__ void dst16_tail(int *s, int *o) {
  // index names: s[10]=sa s[11]=sb s[12]=sc s[13]=sd s[14]=se s[15]=sf
  // Stage 1
  b_sub_a(hf(s[0]), s[7], s[0], s[7]);
  b_sub_a(hf(s[8]), s[15], s[8], s[15]);
  b_add_a(hf(s[4]), s[3], s[4], s[3]);
  b_add_a(hf(s[12]), s[11], s[12], s[11]);
  b_sub_a(hf(s[2]), s[5], s[2], s[5]);
  b_sub_a(hf(s[10]), s[13], s[10], s[13]);
  b_add_a(hf(s[6]), s[1], s[6], s[1]);
  b_add_a(hf(s[14]), s[9], s[14], s[9]);
  // Stage 2
  HF S8, S7, SA, S5, S0, SD, S2, SF;
  int h4, hb, h6, h9, h3, he, h1, hc;
  b_add(s[8], s[4], S8, h4);
  b_add(s[7], s[11], S7, hb);
  b_sub(s[10], s[6], SA, h6);
  b_sub(s[5], s[9], S5, h9);
  b_add(s[0], s[3], S0, h3);
  b_add(s[13], s[14], SD, he);
  b_sub(s[2], s[1], S2, h1);
  b_sub(s[15], s[12], SF, hc);
  int s8 = S8.f, s7 = S7.f, sa = SA.f, s5 = S5.f;  // the full values are carried on
  int s9, s6, s4, sb;
  // Stage 3
  if (!ASYM) {
    rot_f<R_ADDAVG>(s8, s7, 301, 8, 1609, 11, 12785, 15, s8, s7);
    rot_f<R_ADD>(h9, h6, 11363, 13, 9041, 15, 4551, 13, s9, s6);
    rot_n<true>(s5, sa, 5681, 12, 9041, 15, 4551, 12, s5, sa);
    rot_n<false>(h4, hb, 9633, 13, 12873, 14, 6393, 15, s4, sb);
  } else {
    rot_f<R_ADD>(s8, s7, 9633, 13, 12873, 14, 6393, 15, s8, s7);
    rot_f<R_ADD>(h9, h6, 22725, 14, 9041, 15, 4551, 13, s9, s6);
    rot_n<false>(s5, sa, 11363, 13, 9041, 15, 4551, 13, s5, sa);
    rot_n<false>(h4, hb, 9633, 13, 12873, 14, 6393, 15, s4, sb);
  }
  // Stage 4
  int s2, sc, s0, s1, sf, se, sd, s3;
  b_add_a(S2, hc, s2, sc);
  b_sub_a(S0, h1, s0, s1);
  b_add_a(SF, he, sf, se);
  b_add_a(SD, h3, sd, s3);
  b_add_a(hf(s7), s6, s7, s6);
  b_sub_a(hf(s8), s9, s8, s9);
  b_sub_a(hf(sa), sb, sa, sb);
  b_add_a(hf(s5), s4, s5, s4);
  // Stage 5
  if (!ASYM) {
    rot_f<R_ADDAVG>(sc, s3, 669, 9, 8867, 14, 3135, 12, sc, s3);
    rot_n<true>(s2, sd, 669, 9, 8867, 14, 3135, 12, s2, sd);
    rot_q<Q_ADDAVG>(sa, s5, 5793, 12, 11585, 13, sa, s5);
    rot_q<Q_ADDAVG>(s6, s9, 5793, 12, 11585, 13, s6, s9);
    rot_q<Q_ADDAVG>(se, s1, 5793, 12, 11585, 13, se, s1);
  } else {
    rot_f<R_ADD>(sc, s3, 10703, 13, 8867, 14, 3135, 13, sc, s3);
    rot_n<false>(s2, sd, 10703, 13, 8867, 14, 3135, 13, s2, sd);
    rot_q<Q_ADD>(sa, s5, 11585, 13, 5793, 13, sa, s5);
    rot_q<Q_ADD>(s6, s9, 11585, 13, 5793, 13, s6, s9);
    rot_q<Q_ADD>(se, s1, 11585, 13, 5793, 13, se, s1);
  }
  o[0] = s0; o[1] = s1; o[2] = s2; o[3] = s3; o[4] = s4; o[5] = s5; o[6] = s6; o[7] = s7;
  o[8] = s8; o[9] = s9; o[10] = sa; o[11] = sb; o[12] = sc; o[13] = sd; o[14] = se; o[15] = sf;
}
// forward.rs:751-873
__device__ __forceinline__ void dst16_core(const int *x, int *o) {
  int s[16];
  rot_f<R_ADDSHIFT>(x[0], x[15], 24279, 15, 11003, 13, 1137, 14, s[0], s[15]);
  rot_f<R_SUBSHIFT>(x[14], x[1], 1645, 11, 305, 8, 425, 11, s[14], s[1]);
  rot_f<R_ADDSHIFT>(x[2], x[13], 14053, 14, 8423, 13, 2815, 13, s[2], s[13]);
  rot_f<R_SUBSHIFT>(x[12], x[3], 14811, 14, 7005, 13, 3903, 13, s[12], s[3]);
  rot_f<R_ADDSHIFT>(x[4], x[11], 30853, 15, 11039, 14, 9907, 14, s[4], s[11]);
  rot_f<R_SUBSHIFT>(x[10], x[5], 15893, 14, 3981, 13, 1489, 11, s[10], s[5]);
  rot_f<R_ADDSHIFT>(x[6], x[9], 32413, 15, 601, 11, 13803, 14, s[6], s[9]);
  rot_f<R_SUBSHIFT>(x[8], x[7], 32729, 15, 201, 11, 1945, 11, s[8], s[7]);
  dst16_tail<false>(s, o);
}
// forward.rs:938-1070 : inputs (s0, s1h, s2, s3h, ..., se, sfh): P[] even, h[] odd
__device__ __forceinline__ void dst16_asym(const HF *P, const int *h, int *o) {
  int s[16];
  rot_h<R_ADDSHIFT>(P[0], h[15], 1073, 11, 62241, 15, 201, 11, s[0], s[15]);
  rot_h<R_SUBSHIFT>(P[14], h[1], 18611, 15, 55211, 15, 601, 11, s[14], s[1]);
  rot_h<R_ADDSHIFT>(P[2], h[13], 9937, 14, 1489, 10, 3981, 13, s[2], s[13]);
  rot_h<R_SUBSHIFT>(P[12], h[3], 10473, 14, 39627, 15, 11039, 14, s[12], s[3]);
  rot_h<R_ADDSHIFT>(P[4], h[11], 2727, 12, 3903, 12, 7005, 13, s[4], s[11]);
  rot_h<R_SUBSHIFT>(P[10], h[5], 5619, 13, 2815, 12, 8423, 13, s[10], s[5]);
  rot_h<R_ADDSHIFT>(P[6], h[9], 2865, 12, 13599, 15, 305, 8, s[6], s[9]);
  rot_h<R_SUBSHIFT>(P[8], h[7], 23143, 15, 1137, 13, 11003, 13, s[8], s[7]);
  dst16_tail<true>(s, o);
}

// --------------------------------------------------------------- 32-point ----
// forward.rs:1072-1136
__device__ __forceinline__ void dct32_core(const int *t, int *o) {
  int h[16];
  HF P[16];
  dct_split<32>(t, h, P);
  dct16_asym(h, P, o);
  dst16_asym(P, h, o + 16);
  reverse_in_place<16>(o + 16);
}
// forward.rs:1212-1277
__device__ __forceinline__ void dct32_asym(const int *h, const HF *P, int *o) {
  int t[32];
  dct_asym_merge<32>(h, P, t);
  dct16_core(t, o);
  int r[16];
#pragma unroll
  for (int i = 0; i < 16; i++) r[i] = t[31 - i];  // (tv, tu, ..., tg)
  dst16_core(r, o + 16);
  reverse_in_place<16>(o + 16);
}

// forward.rs:1279-1548 : inputs (t0, t1h, t2, t3h, ..., tu, tvh): P[] even, h[] odd.
// Index names: 10=a 11=b 12=c 13=d 14=e 15=f 16=g 17=h 18=i 19=j 20=k 21=l
//              22=m 23=n 24=o 25=p 26=q 27=r 28=s 29=t 30=u 31=v
__device__ __forceinline__ void dst32_asym(const HF *P, const int *h, int *o) {
  int t[32];
  // Stage 0 (:1290-1366): even i RotateAdd::half(t_i, t_(31-i)h) -> (t_i, t_(31-i));
  //                       odd  i RotateSub::half(t_(31-i), t_i h) -> (t_(31-i), t_i)
  rot_h<R_ADD>(P[0], h[31], 5933, 13, 22595, 14, 1137, 15, t[0], t[31]);
  rot_h<R_SUB>(P[30], h[1], 6203, 13, 21403, 14, 3409, 15, t[30], t[1]);
  rot_h<R_ADD>(P[2], h[29], 25833, 15, 315, 8, 5673, 15, t[2], t[29]);
  rot_h<R_SUB>(P[28], h[3], 26791, 15, 4717, 12, 7923, 15, t[28], t[3]);
  rot_h<R_ADD>(P[4], h[27], 6921, 13, 17531, 14, 10153, 15, t[4], t[27]);
  rot_h<R_SUB>(P[26], h[5], 28511, 15, 32303, 15, 1545, 12, t[26], t[5]);
  rot_h<R_ADD>(P[6], h[25], 29269, 15, 14733, 14, 1817, 12, t[6], t[25]);
  rot_h<R_SUB>(P[24], h[7], 29957, 15, 13279, 14, 8339, 14, t[24], t[7]);
  rot_h<R_ADD>(P[8], h[23], 7643, 13, 11793, 14, 18779, 15, t[8], t[23]);
  rot_h<R_SUB>(P[22], h[9], 15557, 14, 20557, 15, 20835, 15, t[22], t[9]);
  rot_h<R_ADD>(P[10], h[21], 31581, 15, 17479, 15, 22841, 15, t[10], t[21]);
  rot_h<R_SUB>(P[20], h[11], 7993, 13, 14359, 15, 3099, 12, t[20], t[11]);
  rot_h<R_ADD>(P[12], h[19], 16143, 14, 2801, 13, 26683, 15, t[12], t[19]);
  rot_h<R_SUB>(P[18], h[13], 16261, 14, 4011, 14, 14255, 14, t[18], t[13]);
  rot_h<R_ADD>(P[14], h[17], 32679, 15, 4821, 15, 30269, 15, t[14], t[17]);
  rot_h<R_SUB>(P[16], h[15], 16379, 14, 201, 12, 15977, 14, t[16], t[15]);

  // Stage 1 (:1369-1384).  F[k] = pair left in slot k, g[k] = half left in slot k.
  HF F[32];
  int g[32];
  b_add(t[0], t[15], F[0], g[15]);
  b_sub(t[31], t[16], F[31], g[16]);
  b_add(t[17], t[30], F[17], g[30]);
  b_sub(t[14], t[1], F[14], g[1]);
  b_add(t[2], t[13], F[2], g[13]);
  b_sub(t[29], t[18], F[29], g[18]);
  b_add(t[19], t[28], F[19], g[28]);
  b_sub(t[12], t[3], F[12], g[3]);
  b_add(t[4], t[11], F[4], g[11]);
  b_sub(t[27], t[20], F[27], g[20]);
  b_add(t[21], t[26], F[21], g[26]);
  b_sub(t[10], t[5], F[10], g[5]);
  b_add(t[6], t[9], F[6], g[9]);
  b_sub(t[25], t[22], F[25], g[22]);
  b_add(t[23], t[24], F[23], g[24]);
  b_sub(t[8], t[7], F[8], g[7]);

  // Stage 2 (:1387-1402)
  b_sub_a(F[0], g[7], t[0], t[7]);
  b_add_a(F[31], g[24], t[31], t[24]);
  b_sub_a(F[25], g[30], t[25], t[30]);
  b_add_a(F[6], g[1], t[6], t[1]);
  b_sub_a(F[2], g[5], t[2], t[5]);
  b_add_a(F[29], g[26], t[29], t[26]);
  b_sub_a(F[27], g[28], t[27], t[28]);
  b_add_a(F[4], g[3], t[4], t[3]);
  b_add_a(F[8], g[16], t[8], t[16]);
  b_sub_a(F[14], g[22], t[14], t[22]);
  b_add_a(F[23], g[15], t[23], t[15]);
  b_sub_a(F[17], g[9], t[17], t[9]);
  b_add_a(F[10], g[18], t[10], t[18]);
  b_sub_a(F[12], g[20], t[12], t[20]);
  b_add_a(F[21], g[13], t[21], t[13]);
  b_sub_a(F[19], g[11], t[19], t[11]);

  // Stage 3 (:1408-1444)
  rot_f<R_SUB>(t[15], t[16], 17911, 14, 14699, 14, 803, 13, t[15], t[16]);
  rot_f<R_ADD>(t[17], t[14], 10217, 13, 5461, 13, 1189, 12, t[17], t[14]);
  rot_f<R_ADD>(t[18], t[13], 5543, 12, 3363, 13, 7723, 14, t[18], t[13]);
  rot_f<R_SUB>(t[12], t[19], 11529, 13, 2271, 14, 5197, 13, t[12], t[19]);
  rot_n<false>(t[11], t[20], 11529, 13, 2271, 14, 5197, 13, t[11], t[20]);
  rot_n<false>(t[10], t[21], 5543, 12, 3363, 13, 7723, 14, t[10], t[21]);
  rot_n<false>(t[9], t[22], 10217, 13, 5461, 13, 1189, 12, t[9], t[22]);
  rot_n<false>(t[8], t[23], 17911, 14, 14699, 14, 803, 13, t[8], t[23]);

  // Stage 4 (:1447-1462)
  b_sub(t[3], t[0], F[3], g[0]);
  b_add(t[28], t[31], F[28], g[31]);
  b_sub(t[30], t[29], F[30], g[29]);
  b_add(t[1], t[2], F[1], g[2]);
  b_add(t[24], t[4], F[24], g[4]);   // ((_toh, to), t4h)
  b_sub(t[26], t[6], F[26], g[6]);   // ((_tqh, tq), t6h)
  b_add(t[7], t[27], F[7], g[27]);   // ((_t7h, t7), trh)
  b_sub(t[5], t[25], F[5], g[25]);   // ((_t5h, t5), tph)
  b_sub(t[11], t[8], F[11], g[8]);
  b_add(t[20], t[23], F[20], g[23]);
  b_sub(t[22], t[21], F[22], g[21]);
  b_add(t[9], t[10], F[9], g[10]);
  b_sub(t[15], t[12], F[15], g[12]);
  b_add(t[16], t[19], F[16], g[19]);
  b_sub(t[18], t[17], F[18], g[17]);
  b_add(t[13], t[14], F[13], g[14]);

  // Stage 5 (:1468-1483): only the full values of to, tq, t7, t5 are used
  int to_ = F[24].f, tq_ = F[26].f, t7_ = F[7].f, t5_ = F[5].f;
  rot_f<R_ADD>(to_, t7_, 301, 8, 1609, 11, 6393, 15, to_, t7_);
  rot_f<R_ADD>(g[25], g[6], 11363, 13, 9041, 15, 4551, 13, g[25], g[6]);   // (tph, t6h)
  rot_n<false>(t5_, tq_, 5681, 12, 9041, 15, 4551, 13, t5_, tq_);
  rot_n<false>(g[4], g[27], 9633, 13, 12873, 14, 6393, 15, g[4], g[27]);   // (t4h, trh)

  // Stage 6 (:1486-1501)
  b_add_a(F[1], g[0], t[1], t[0]);
  b_sub_a(F[30], g[31], t[30], t[31]);
  b_sub_a(F[28], g[2], t[28], t[2]);
  b_sub_a(F[3], g[29], t[3], t[29]);
  b_add_a(hf(t5_), g[4], t[5], t[4]);
  b_sub_a(hf(tq_), g[27], t[26], t[27]);
  b_add_a(hf(t7_), g[6], t[7], t[6]);
  b_sub_a(hf(to_), g[25], t[24], t[25]);
  b_add_a(F[9], g[8], t[9], t[8]);
  b_sub_a(F[22], g[23], t[22], t[23]);
  b_sub_a(F[20], g[10], t[20], t[10]);
  b_sub_a(F[11], g[21], t[11], t[21]);
  b_add_a(F[18], g[12], t[18], t[12]);
  b_add_a(F[13], g[19], t[13], t[19]);
  b_add_a(F[15], g[14], t[15], t[14]);
  b_sub_a(F[16], g[17], t[16], t[17]);

  // Stage 7 (:1507-1542)
  rot_n<false>(t[2], t[29], 669, 9, 8867, 14, 3135, 13, t[2], t[29]);
  rot_f<R_ADD>(t[28], t[3], 669, 9, 8867, 14, 3135, 13, t[28], t[3]);
  rot_n<false>(t[10], t[21], 669, 9, 8867, 14, 3135, 13, t[10], t[21]);
  rot_f<R_ADD>(t[20], t[11], 669, 9, 8867, 14, 3135, 13, t[20], t[11]);
  rot_f<R_ADD>(t[12], t[19], 669, 9, 8867, 14, 3135, 13, t[12], t[19]);
  rot_n<false>(t[18], t[13], 669, 9, 8867, 14, 3135, 13, t[18], t[13]);
  rot_q<Q_ADD>(t[30], t[1], 5793, 12, 5793, 13, t[30], t[1]);
  rot_q<Q_ADD>(t[26], t[5], 5793, 12, 5793, 13, t[26], t[5]);
  rot_q<Q_SUB>(t[25], t[6], 5793, 12, 5793, 13, t[25], t[6]);
  rot_q<Q_ADD>(t[22], t[9], 5793, 12, 5793, 13, t[22], t[9]);
  rot_q<Q_ADD>(t[14], t[17], 5793, 12, 5793, 13, t[14], t[17]);
#pragma unroll
  for (int i = 0; i < 32; i++) o[i] = t[i];
}

// --------------------------------------------------------------- 64-point ----
// forward.rs:1551-1684 (without the final reorder; see perm below)
__device__ __forceinline__ void dct64_core(const int *x, int *o) {
  // butterfly_pair(i): j = 2i: (ah, c) = neg(x[j], x[63-j]); (b, dh) = add(x[j+1], x[62-j]);
  // half[i] = ah, half[31-i] = dh, asym[i] = b, asym[31-i] = c
  int half_[32];
  HF asym[32];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const int j = 2 * i;
    b_neg(x[j], x[63 - j], half_[i], asym[31 - i]);
    b_add(x[j + 1], x[62 - j], asym[i], half_[31 - i]);
  }
  int h[32];
  HF P[32];
  // (half[0], asym[0], half[1], asym[1], ..., half[15], asym[15])
#pragma unroll
  for (int k = 0; k < 16; k++) {
    h[2 * k] = half_[k];
    P[2 * k + 1] = asym[k];
  }
  dct32_asym(h, P, o);
  // (asym[31], half[31], asym[30], half[30], ..., asym[16], half[16])
#pragma unroll
  for (int k = 0; k < 16; k++) {
    P[2 * k] = asym[31 - k];
    h[2 * k + 1] = half_[31 - k];
  }
  dst32_asym(P, h, o + 32);
  reverse_in_place<32>(o + 32);
}

// ------------------------------------------------------------ entry points ----
// All even-size kernels leave their result in bit-reversed order; the
// reference undoes it with explicit index lists (forward.rs:369-372,499-506,
// 733-748,1178-1209,1661-1683) which are exactly the log2(N)-bit reversal.
template <int N>
__host__ __device__ constexpr int bitrev(int i) {
  int r = 0;
  for (int b = 1, n = N >> 1; n; b <<= 1, n >>= 1)
    if (i & b) r |= n;
  return r;
}

enum Tx1D { TX1_DCT = 0, TX1_ADST = 1, TX1_ID = 2 };

// in[] natural order -> out[] natural frequency order.  T = Tx1D.
template <int T, int N>
__device__ __forceinline__ void fwd1d(const int *in, int *out) {
  if constexpr (T == TX1_ID) {
#pragma unroll
    for (int i = 0; i < N; i++) out[i] = in[i];
  } else if constexpr (T == TX1_ADST && N == 4) {
    // daala_fdst_vii_4, forward.rs:375-405 (already in natural order)
    const int q0 = in[0], q1 = in[1], q2 = in[2], q3 = in[3];
    int t0 = wadd(q1, q3);
    int t1 = wadd(q1, avg_m(q0, t0));
    int t2 = wsub(q0, q1);
    int t3 = q2;
    int t4 = wadd(q0, q3);
    t0 = mulc(t0, 7021, 14);
    t1 = mulc(t1, 37837, 15);
    t2 = mulc(t2, 21513, 15);
    t3 = mulc(t3, 37837, 15);
    t4 = mulc(t4, 467, 11);
    const int t3h = half0(t3);
    const int u4 = wadd(t4, t3h);
    out[0] = wadd(t0, u4);
    out[1] = t1;
    out[2] = wadd(t0, wsub(t2, t3h));
    out[3] = wadd(t2, wsub(t3, u4));
  } else {
    int t[N];
    if constexpr (T == TX1_DCT) {
      if constexpr (N == 4) dct4_core(in[0], in[1], in[2], in[3], t);
      else if constexpr (N == 8) dct8_core(in, t);
      else if constexpr (N == 16) dct16_core(in, t);
      else if constexpr (N == 32) dct32_core(in, t);
      else dct64_core(in, t);
    } else {
      static_assert(T != TX1_ADST || N == 8 || N == 16, "ADST exists for 4, 8, 16 only");
      if constexpr (N == 8) dst8_core(in, t);
      else dst16_core(in, t);
    }
#pragma unroll
    for (int i = 0; i < N; i++) out[i] = t[bitrev<N>(i)];
  }
}

}  // namespace tx
