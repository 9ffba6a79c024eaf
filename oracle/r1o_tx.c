/*
 * r1o_tx.c -- ORACLE (test infrastructure, never shipped): forward
 * transforms restated from src/transform/forward.rs (scalar `impl
 * TxOperations for i32`, forward.rs:114-130) and the 2-D driver
 * FwdTxfm2D::fht (forward.rs:1804-1899).
 *
 * Parity: PINNED against the reference's own functions.  The reference holds
 * no golden coefficients (only a fwd->inv round-trip tolerance test,
 * src/transform/mod.rs:589-725), but tests/ref_fwd_tx.py runs the text of
 * forward.rs:100-1700 itself (transliterated at test time, dev container
 * only): tests/test_oracle_vs_reference_src.py compares every 1-D kernel and
 * all 193 2-D pairs with it, and tests/golden/fwd_tx_r01.npz holds vectors
 * made the same way.  The round-trip test is run too (test_oracle_inv_tx.py).
 * Constants are taken from the CODE, never from the comments (several differ,
 * SURVEY.md A1).  tests/test_oracle_props.py also compares every 1-D kernel
 * against the floating-point DCT-II / DST-IV / DST-VII it approximates.
 *
 * Arithmetic is 32-bit two's complement with wrap-around (release-mode Rust
 * and every SIMD lane wrap): build with -fwrapv.  `>>` on negative values is
 * an arithmetic shift (gcc guarantees this).
 *
 * Naming: a lower-case variable is a plain value; an upper-case one is the
 * reference's asymmetric pair (half, full) -- `(T, T)` in the Rust.
 */
#include "r1o.h"
#include <string.h>

typedef int32_t T;
typedef struct { T h, f; } P; /* (half, full) */

/* forward.rs:115-129 */
static inline T mul(T x, int32_t c, int s) { return (T)((x * c) + ((1 << s) >> 1)) >> s; }
static inline T rs1(T x) { return (x + (x < 0 ? 1 : 0)) >> 1; }
static inline T avg_add(T a, T b) { return (a + b) >> 1; }
static inline T avg_sub(T a, T b) { return (a - b) >> 1; }
static inline P mkp(T h, T f) { P p = {h, f}; return p; }
/* (x.rshift1(), x) */
static inline P hp(T x) { return mkp(rs1(x), x); }

/* ---- rotation kernels ---------------------------------------------------- */
enum { K_ADD, K_ADDAVG, K_ADDSHIFT, K_SUB, K_SUBAVG, K_SUBSHIFT };

/* RotateKernel::half_kernel, forward.rs:206-215 with the six impls :243-277 */
static inline void rot_half(int k, P p0, T p1, int c0, int s0, int c1, int s1, int c2, int s2,
                            T *o0, T *o1) {
  T t;
  switch (k) {
    case K_ADD: case K_ADDSHIFT: t = p1 + p0.h; break;
    case K_ADDAVG: t = avg_add(p1, p0.h); break;
    case K_SUB: case K_SUBSHIFT: t = p1 - p0.h; break;
    default: t = avg_sub(p1, p0.h); break;
  }
  T a = mul(p0.f, c0, s0), b = mul(p1, c1, s1), c = mul(t, c2, s2);
  T sh = (k == K_ADDSHIFT || k == K_SUBSHIFT) ? rs1(c) : c;
  *o0 = b + c;
  *o1 = (k == K_ADD || k == K_ADDAVG || k == K_ADDSHIFT) ? a - sh : a + sh;
}
/* RotateKernel::kernel, forward.rs:217-219 */
static inline void rot(int k, T p0, T p1, int c0, int s0, int c1, int s1, int c2, int s2, T *o0,
                       T *o1) {
  rot_half(k, mkp(p0, p0), p1, c0, s0, c1, s1, c2, s2, o0, o1);
}
/* RotateKernelNeg::kernel, forward.rs:222-232; avg=0 RotateNeg, 1 RotateNegAvg */
static inline void rot_neg(int avg, T p0, T p1, int c0, int s0, int c1, int s1, int c2, int s2,
                           T *o0, T *o1) {
  T t = avg ? avg_sub(p0, p1) : p0 - p1;
  T a = mul(p0, c0, s0), b = mul(p1, c1, s1), c = mul(t, c2, s2);
  *o0 = b - c;
  *o1 = c - a;
}
/* RotateKernelPi4::kernel, forward.rs:164-199 */
enum { PI4_ADD, PI4_ADDAVG, PI4_SUB, PI4_SUBAVG };
static inline void rot_pi4(int k, T p0, T p1, int c0, int s0, int c1, int s1, T *o0, T *o1) {
  T t;
  switch (k) {
    case PI4_ADD: t = p1 + p0; break;
    case PI4_ADDAVG: t = avg_add(p1, p0); break;
    case PI4_SUB: t = p1 - p0; break;
    default: t = avg_sub(p1, p0); break;
  }
  T a = mul(p0, c0, s0), r0 = mul(t, c1, s1);
  *o0 = r0;
  *o1 = (k == PI4_ADD || k == PI4_ADDAVG) ? a - r0 : a + r0;
}

/* ---- butterflies, forward.rs:287-324 -------------------------------------- */
static inline void bf_add(T p0, T p1, P *o0, T *o1h) {
  T s = p0 + p1, h = rs1(s);
  *o0 = mkp(h, s);
  *o1h = p1 - h;
}
static inline void bf_sub(T p0, T p1, P *o0, T *o1h) {
  T s = p0 - p1, h = rs1(s);
  *o0 = mkp(h, s);
  *o1h = p1 + h;
}
static inline void bf_neg(T p0, T p1, T *o0h, P *o1) {
  T d = p0 - p1, h = rs1(d);
  *o0h = p0 - h;
  *o1 = mkp(h, d);
}
static inline void bf_add_asym(P p0, T p1h, T *o0, T *o1) {
  T p1 = p1h + p0.h;
  *o0 = p0.f - p1;
  *o1 = p1;
}
static inline void bf_sub_asym(P p0, T p1h, T *o0, T *o1) {
  T p1 = p1h - p0.h;
  *o0 = p0.f + p1;
  *o1 = p1;
}
static inline void bf_neg_asym(T p0h, P p1, T *o0, T *o1) {
  T p0 = p0h + p1.h;
  *o0 = p0;
  *o1 = p0 - p1.f;
}

static inline void rev(T *a, int n) {
  for (int i = 0; i < n / 2; i++) { T t = a[i]; a[i] = a[n - 1 - i]; a[n - 1 - i] = t; }
}

/* ---- 2-point ---------------------------------------------------------------- */
/* forward.rs:338-340 */
static inline void fdct_ii_2_asym(T p0h, P p1, T *o0, T *o1) { bf_neg_asym(p0h, p1, o0, o1); }
/* forward.rs:342-347 */
static inline void fdst_iv_2_asym(P p0, T p1h, T *o0, T *o1) {
  rot_half(K_ADD, p0, p1h, 473, 9, 3135, 12, 4433, 13, o0, o1);
}
/* forward.rs:407-412: (p1,p0) = Pi4SubAvg(p1,p0) ; return (p0,p1) */
static inline void fdct_ii_2(T p0, T p1, T *o0, T *o1) {
  T n1, n0;
  rot_pi4(PI4_SUBAVG, p1, p0, 11585, 13, 11585, 13, &n1, &n0);
  *o0 = n0;
  *o1 = n1;
}
/* forward.rs:414-419 */
static inline void fdst_iv_2(T p0, T p1, T *o0, T *o1) {
  rot(K_ADDAVG, p0, p1, 10703, 13, 8867, 14, 3135, 12, o0, o1);
}

/* ---- 4-point ---------------------------------------------------------------- */
/* forward.rs:349-361 */
static void fdct_ii_4(T q0, T q1, T q2, T q3, T *out) {
  T q0h, q2h; P Q3, Q1;
  bf_neg(q0, q3, &q0h, &Q3);
  bf_add(q1, q2, &Q1, &q2h);
  fdct_ii_2_asym(q0h, Q1, &q0, &q1);
  fdst_iv_2_asym(Q3, q2h, &q3, &q2);
  out[0] = q0; out[1] = q1; out[2] = q2; out[3] = q3;
}
/* forward.rs:363-373 */
static void fdct4(const T *in, T *out) {
  T t[4];
  fdct_ii_4(in[0], in[1], in[2], in[3], t);
  out[0] = t[0]; out[1] = t[2]; out[2] = t[1]; out[3] = t[3];
}
/* forward.rs:375-405 */
static void fdst_vii_4(const T *in, T *out) {
  T q0 = in[0], q1 = in[1], q2 = in[2], q3 = in[3];
  T t0 = q1 + q3;
  T t1 = q1 + avg_sub(q0, t0);
  T t2 = q0 - q1;
  T t3 = q2;
  T t4 = q0 + q3;
  t0 = mul(t0, 7021, 14);
  t1 = mul(t1, 37837, 15);
  t2 = mul(t2, 21513, 15);
  t3 = mul(t3, 37837, 15);
  t4 = mul(t4, 467, 11);
  T t3h = rs1(t3);
  T u4 = t4 + t3h;
  out[0] = t0 + u4;
  out[1] = t1;
  out[2] = t0 + (t2 - t3h);
  out[3] = t2 + (t3 - u4);
}
/* forward.rs:421-433 */
static void fdct_ii_4_asym(T q0h, P Q1, T q2h, P Q3, T *out) {
  T q0, q1, q2, q3;
  bf_neg_asym(q0h, Q3, &q0, &q3);
  bf_sub_asym(Q1, q2h, &q1, &q2);
  fdct_ii_2(q0, q1, &q0, &q1);
  fdst_iv_2(q3, q2, &q3, &q2);
  out[0] = q0; out[1] = q1; out[2] = q2; out[3] = q3;
}
/* forward.rs:435-466 */
static void fdst_iv_4_asym(P Q0, T q1h, P Q2, T q3h, T *out) {
  T q0, q1, q2, q3;
  rot_half(K_ADDSHIFT, Q0, q3h, 9633, 14, 12873, 13, 12785, 15, &q0, &q3);
  rot_half(K_SUBSHIFT, Q2, q1h, 11363, 14, 18081, 15, 4551, 12, &q2, &q1);
  bf_sub_asym(hp(q2), q3, &q2, &q3);
  bf_sub_asym(hp(q0), q1, &q0, &q1);
  rot_pi4(PI4_ADDAVG, q2, q1, 11585, 13, 11585, 13, &q2, &q1);
  out[0] = q0; out[1] = q1; out[2] = q2; out[3] = q3;
}
/* forward.rs:590-615 */
static void fdst_iv_4(T q0, T q1, T q2, T q3, T *out) {
  rot(K_ADDSHIFT, q0, q3, 13623, 14, 4551, 12, 565, 11, &q0, &q3);
  rot(K_SUBSHIFT, q2, q1, 16069, 14, 12785, 15, 1609, 11, &q2, &q1);
  bf_sub_asym(hp(q2), q3, &q2, &q3);
  bf_sub_asym(hp(q0), q1, &q0, &q1);
  rot_pi4(PI4_ADDAVG, q2, q1, 11585, 13, 11585, 13, &q2, &q1);
  out[0] = q0; out[1] = q1; out[2] = q2; out[3] = q3;
}

/* ---- 8-point ---------------------------------------------------------------- */
/* forward.rs:468-481 */
static void fdct_ii_8(T r0, T r1, T r2, T r3, T r4, T r5, T r6, T r7, T *out) {
  T r0h, r6h, r2h, r4h; P R7, R1, R5, R3;
  bf_neg(r0, r7, &r0h, &R7);
  bf_add(r1, r6, &R1, &r6h);
  bf_neg(r2, r5, &r2h, &R5);
  bf_add(r3, r4, &R3, &r4h);
  fdct_ii_4_asym(r0h, R1, r2h, R3, out);
  fdst_iv_4_asym(R7, r6h, R5, r4h, out + 4);
  rev(out + 4, 4);
}
static const uint8_t PERM8[8] = {0, 4, 2, 6, 1, 5, 3, 7};
/* forward.rs:483-507 */
static void fdct8(const T *in, T *out) {
  T t[8];
  fdct_ii_8(in[0], in[1], in[2], in[3], in[4], in[5], in[6], in[7], t);
  for (int i = 0; i < 8; i++) out[i] = t[PERM8[i]];
}
/* forward.rs:509-562 */
static void fdst_iv_8(T r0, T r1, T r2, T r3, T r4, T r5, T r6, T r7, T *out) {
  rot(K_ADD, r0, r7, 17911, 14, 14699, 14, 803, 13, &r0, &r7);
  rot(K_SUB, r6, r1, 20435, 14, 21845, 15, 1189, 12, &r6, &r1);
  rot(K_ADD, r2, r5, 22173, 14, 3363, 13, 15447, 15, &r2, &r5);
  rot(K_SUB, r4, r3, 23059, 14, 2271, 14, 5197, 13, &r4, &r3);
  P R0, R2, R5, R7; T r3h, r1h, r6h, r4h;
  bf_add(r0, r3, &R0, &r3h);
  bf_sub(r2, r1, &R2, &r1h);
  bf_add(r5, r6, &R5, &r6h);
  bf_sub(r7, r4, &R7, &r4h);
  bf_add_asym(R7, r6h, &r7, &r6);
  bf_add_asym(R5, r3h, &r5, &r3);
  bf_add_asym(R2, r4h, &r2, &r4);
  bf_sub_asym(R0, r1h, &r0, &r1);
  rot(K_SUBAVG, r3, r4, 10703, 13, 8867, 14, 3135, 12, &r3, &r4);
  rot_neg(1, r2, r5, 10703, 13, 8867, 14, 3135, 12, &r2, &r5);
  rot_pi4(PI4_SUBAVG, r1, r6, 11585, 13, 11585, 13, &r1, &r6);
  out[0] = r0; out[1] = r1; out[2] = r2; out[3] = r3;
  out[4] = r4; out[5] = r5; out[6] = r6; out[7] = r7;
}
/* forward.rs:564-588 */
static void fdst8(const T *in, T *out) {
  T t[8];
  fdst_iv_8(in[0], in[1], in[2], in[3], in[4], in[5], in[6], in[7], t);
  for (int i = 0; i < 8; i++) out[i] = t[PERM8[i]];
}
/* forward.rs:617-631 */
static void fdct_ii_8_asym(T r0h, P R1, T r2h, P R3, T r4h, P R5, T r6h, P R7, T *out) {
  T r0, r1, r2, r3, r4, r5, r6, r7;
  bf_neg_asym(r0h, R7, &r0, &r7);
  bf_sub_asym(R1, r6h, &r1, &r6);
  bf_neg_asym(r2h, R5, &r2, &r5);
  bf_sub_asym(R3, r4h, &r3, &r4);
  fdct_ii_4(r0, r1, r2, r3, out);
  fdst_iv_4(r7, r6, r5, r4, out + 4);
  rev(out + 4, 4);
}
/* forward.rs:633-687 */
static void fdst_iv_8_asym(P R0, T r1h, P R2, T r3h, P R4, T r5h, P R6, T r7h, T *out) {
  T r0, r1, r2, r3, r4, r5, r6, r7;
  rot_half(K_ADD, R0, r7h, 12665, 14, 5197, 12, 2271, 14, &r0, &r7);
  rot_half(K_SUB, R6, r1h, 14449, 14, 30893, 15, 3363, 13, &r6, &r1);
  rot_half(K_ADD, R2, r5h, 15679, 14, 1189, 11, 5461, 13, &r2, &r5);
  rot_half(K_SUB, R4, r3h, 16305, 14, 803, 12, 14699, 14, &r4, &r3);
  P S0, S2, S5, S7; T s3h, s1h, s6h, s4h;
  bf_add(r0, r3, &S0, &s3h);
  bf_sub(r2, r1, &S2, &s1h);
  bf_add(r5, r6, &S5, &s6h);
  bf_sub(r7, r4, &S7, &s4h);
  bf_add_asym(S7, s6h, &r7, &r6);
  bf_add_asym(S5, s3h, &r5, &r3);
  bf_add_asym(S2, s4h, &r2, &r4);
  bf_sub_asym(S0, s1h, &r0, &r1);
  rot(K_SUBAVG, r3, r4, 669, 9, 8867, 14, 3135, 12, &r3, &r4);
  rot_neg(1, r2, r5, 669, 9, 8867, 14, 3135, 12, &r2, &r5);
  rot_pi4(PI4_SUBAVG, r1, r6, 5793, 12, 11585, 13, &r1, &r6);
  out[0] = r0; out[1] = r1; out[2] = r2; out[3] = r3;
  out[4] = r4; out[5] = r5; out[6] = r6; out[7] = r7;
}

/* ---- 16-point --------------------------------------------------------------- */
/* forward.rs:689-707.  s[] = 16 inputs */
static void fdct_ii_16(const T *s, T *out) {
  T h[8]; P A[8]; /* even slots -> half + pair from bf_neg, odd -> pair + half from bf_add */
  /* (s0h,sf)=neg(s0,sf) (s1,seh)=add(s1,se) (s2h,sd)=neg(s2,sd) ... (s7,s8h)=add(s7,s8) */
  for (int i = 0; i < 8; i += 2) {
    bf_neg(s[i], s[15 - i], &h[i], &A[i]);         /* h[i]=s_i h ; A[i]= pair of s_(15-i) */
    bf_add(s[i + 1], s[14 - i], &A[i + 1], &h[i + 1]); /* A[i+1]= pair of s_(i+1) ; h[i+1]= s_(14-i) h */
  }
  /* fdct_ii_8_asym(s0h, s1, s2h, s3, s4h, s5, s6h, s7) */
  fdct_ii_8_asym(h[0], A[1], h[2], A[3], h[4], A[5], h[6], A[7], out);
  /* fdst_iv_8_asym(sf, seh, sd, sch, sb, sah, s9, s8h) */
  fdst_iv_8_asym(A[0], h[1], A[2], h[3], A[4], h[5], A[6], h[7], out + 8);
  rev(out + 8, 8);
}
static const uint8_t PERM16[16] = {0, 8, 4, 12, 2, 10, 6, 14, 1, 9, 5, 13, 3, 11, 7, 15};
/* forward.rs:709-749 */
static void fdct16(const T *in, T *out) {
  T t[16];
  fdct_ii_16(in, t);
  for (int i = 0; i < 16; i++) out[i] = t[PERM16[i]];
}

/* Stages 1..5 of the 16-point DST-IV are shared in shape between
 * daala_fdst_iv_16 (forward.rs:797-872) and daala_fdst_iv_16_asym
 * (forward.rs:994-1069); they differ only in the stage-3 and stage-5 rotation
 * kernels/constants, selected by `asym`.  s[] holds s0..sf after stage 0. */
static void fdst_iv_16_tail(T *s, int asym, T *out) {
  T s0 = s[0], s1 = s[1], s2 = s[2], s3 = s[3], s4 = s[4], s5 = s[5], s6 = s[6], s7 = s[7];
  T s8 = s[8], s9 = s[9], sa = s[10], sb = s[11], sc = s[12], sd = s[13], se = s[14], sf = s[15];
  /* Stage 1 */
  bf_sub_asym(hp(s0), s7, &s0, &s7);
  bf_sub_asym(hp(s8), sf, &s8, &sf);
  bf_add_asym(hp(s4), s3, &s4, &s3);
  bf_add_asym(hp(sc), sb, &sc, &sb);
  bf_sub_asym(hp(s2), s5, &s2, &s5);
  bf_sub_asym(hp(sa), sd, &sa, &sd);
  bf_add_asym(hp(s6), s1, &s6, &s1);
  bf_add_asym(hp(se), s9, &se, &s9);
  /* Stage 2 */
  P S8, S7, SA, S5, S0, SD, S2, SF;
  T s4h, sbh, s6h, s9h, s3h, seh, s1h, sch;
  bf_add(s8, s4, &S8, &s4h);
  bf_add(s7, sb, &S7, &sbh);
  bf_sub(sa, s6, &SA, &s6h);
  bf_sub(s5, s9, &S5, &s9h);
  bf_add(s0, s3, &S0, &s3h);
  bf_add(sd, se, &SD, &seh);
  bf_sub(s2, s1, &S2, &s1h);
  bf_sub(sf, sc, &SF, &sch);
  s8 = S8.f; s7 = S7.f; sa = SA.f; s5 = S5.f; /* ((_h, full), xh): the full value is kept */
  /* Stage 3 */
  if (!asym) { /* forward.rs:821-837 */
    rot(K_ADDAVG, s8, s7, 301, 8, 1609, 11, 12785, 15, &s8, &s7);
    rot(K_ADD, s9h, s6h, 11363, 13, 9041, 15, 4551, 13, &s9, &s6);
    rot_neg(1, s5, sa, 5681, 12, 9041, 15, 4551, 12, &s5, &sa);
    rot_neg(0, s4h, sbh, 9633, 13, 12873, 14, 6393, 15, &s4, &sb);
  } else { /* forward.rs:1018-1034 */
    rot(K_ADD, s8, s7, 9633, 13, 12873, 14, 6393, 15, &s8, &s7);
    rot(K_ADD, s9h, s6h, 22725, 14, 9041, 15, 4551, 13, &s9, &s6);
    rot_neg(0, s5, sa, 11363, 13, 9041, 15, 4551, 13, &s5, &sa);
    rot_neg(0, s4h, sbh, 9633, 13, 12873, 14, 6393, 15, &s4, &sb);
  }
  /* Stage 4 */
  bf_add_asym(S2, sch, &s2, &sc);
  bf_sub_asym(S0, s1h, &s0, &s1);
  bf_add_asym(SF, seh, &sf, &se);
  bf_add_asym(SD, s3h, &sd, &s3);
  bf_add_asym(hp(s7), s6, &s7, &s6);
  bf_sub_asym(hp(s8), s9, &s8, &s9);
  bf_sub_asym(hp(sa), sb, &sa, &sb);
  bf_add_asym(hp(s5), s4, &s5, &s4);
  /* Stage 5 */
  if (!asym) { /* forward.rs:853-868 */
    rot(K_ADDAVG, sc, s3, 669, 9, 8867, 14, 3135, 12, &sc, &s3);
    rot_neg(1, s2, sd, 669, 9, 8867, 14, 3135, 12, &s2, &sd);
    rot_pi4(PI4_ADDAVG, sa, s5, 5793, 12, 11585, 13, &sa, &s5);
    rot_pi4(PI4_ADDAVG, s6, s9, 5793, 12, 11585, 13, &s6, &s9);
    rot_pi4(PI4_ADDAVG, se, s1, 5793, 12, 11585, 13, &se, &s1);
  } else { /* forward.rs:1050-1065 */
    rot(K_ADD, sc, s3, 10703, 13, 8867, 14, 3135, 13, &sc, &s3);
    rot_neg(0, s2, sd, 10703, 13, 8867, 14, 3135, 13, &s2, &sd);
    rot_pi4(PI4_ADD, sa, s5, 11585, 13, 5793, 13, &sa, &s5);
    rot_pi4(PI4_ADD, s6, s9, 11585, 13, 5793, 13, &s6, &s9);
    rot_pi4(PI4_ADD, se, s1, 11585, 13, 5793, 13, &se, &s1);
  }
  out[0] = s0; out[1] = s1; out[2] = s2; out[3] = s3; out[4] = s4; out[5] = s5;
  out[6] = s6; out[7] = s7; out[8] = s8; out[9] = s9; out[10] = sa; out[11] = sb;
  out[12] = sc; out[13] = sd; out[14] = se; out[15] = sf;
}

/* forward.rs:751-873 */
static void fdst_iv_16(const T *in, T *out) {
  T s[16];
  /* Stage 0 (forward.rs:759-795): pairs (s_i, s_(15-i)); even i AddShift(s_i, s_(15-i)),
   * odd i SubShift(s_(15-i), s_i) */
  rot(K_ADDSHIFT, in[0], in[15], 24279, 15, 11003, 13, 1137, 14, &s[0], &s[15]);
  rot(K_SUBSHIFT, in[14], in[1], 1645, 11, 305, 8, 425, 11, &s[14], &s[1]);
  rot(K_ADDSHIFT, in[2], in[13], 14053, 14, 8423, 13, 2815, 13, &s[2], &s[13]);
  rot(K_SUBSHIFT, in[12], in[3], 14811, 14, 7005, 13, 3903, 13, &s[12], &s[3]);
  rot(K_ADDSHIFT, in[4], in[11], 30853, 15, 11039, 14, 9907, 14, &s[4], &s[11]);
  rot(K_SUBSHIFT, in[10], in[5], 15893, 14, 3981, 13, 1489, 11, &s[10], &s[5]);
  rot(K_ADDSHIFT, in[6], in[9], 32413, 15, 601, 11, 13803, 14, &s[6], &s[9]);
  rot(K_SUBSHIFT, in[8], in[7], 32729, 15, 201, 11, 1945, 11, &s[8], &s[7]);
  fdst_iv_16_tail(s, 0, out);
}
/* forward.rs:875-915 */
static void fdst16(const T *in, T *out) {
  T t[16];
  fdst_iv_16(in, t);
  for (int i = 0; i < 16; i++) out[i] = t[PERM16[i]];
}

/* forward.rs:917-936.  h[i] = half inputs at even i, A[i] = pairs at odd i
 * (s0h, s1, s2h, s3, ... seh, sf). */
static void fdct_ii_16_asym(const T *h, const P *A, T *out) {
  T s[16];
  for (int i = 0; i < 8; i += 2) {
    bf_neg_asym(h[i], A[15 - i], &s[i], &s[15 - i]);         /* (s0,sf)=neg_asym(s0h,sf) */
    bf_sub_asym(A[i + 1], h[14 - i], &s[i + 1], &s[14 - i]); /* (s1,se)=sub_asym(s1,seh) */
  }
  fdct_ii_8(s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7], out);
  fdst_iv_8(s[15], s[14], s[13], s[12], s[11], s[10], s[9], s[8], out + 8);
  rev(out + 8, 8);
}

/* forward.rs:938-1070.  A[i] = pairs at even i, h[i] = halves at odd i
 * (s0, s1h, s2, s3h, ..., se, sfh). */
static void fdst_iv_16_asym(const P *A, const T *h, T *out) {
  T s[16];
  rot_half(K_ADDSHIFT, A[0], h[15], 1073, 11, 62241, 15, 201, 11, &s[0], &s[15]);
  rot_half(K_SUBSHIFT, A[14], h[1], 18611, 15, 55211, 15, 601, 11, &s[14], &s[1]);
  rot_half(K_ADDSHIFT, A[2], h[13], 9937, 14, 1489, 10, 3981, 13, &s[2], &s[13]);
  rot_half(K_SUBSHIFT, A[12], h[3], 10473, 14, 39627, 15, 11039, 14, &s[12], &s[3]);
  rot_half(K_ADDSHIFT, A[4], h[11], 2727, 12, 3903, 12, 7005, 13, &s[4], &s[11]);
  rot_half(K_SUBSHIFT, A[10], h[5], 5619, 13, 2815, 12, 8423, 13, &s[10], &s[5]);
  rot_half(K_ADDSHIFT, A[6], h[9], 2865, 12, 13599, 15, 305, 8, &s[6], &s[9]);
  rot_half(K_SUBSHIFT, A[8], h[7], 23143, 15, 1137, 13, 11003, 13, &s[8], &s[7]);
  fdst_iv_16_tail(s, 1, out);
}

/* ---- 32-point --------------------------------------------------------------- */
/* forward.rs:1072-1136 */
static void fdct_ii_32(const T *t, T *out) {
  T h[16]; P A[16];
  for (int i = 0; i < 16; i += 2) {
    bf_neg(t[i], t[31 - i], &h[i], &A[i]);              /* (t0h,tv)=neg(t0,tv) ... */
    bf_add(t[i + 1], t[30 - i], &A[i + 1], &h[i + 1]);  /* (t1,tuh)=add(t1,tu) ... */
  }
  fdct_ii_16_asym(h, A, out);      /* (t0h, t1, t2h, t3, ..., teh, tf) */
  fdst_iv_16_asym(A, h, out + 16); /* (tv, tuh, tt, tsh, ..., th, tgh) */
  rev(out + 16, 16);
}
static const uint8_t PERM32[32] = {0, 16, 8,  24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                   1, 17, 9,  25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};
/* forward.rs:1138-1210 */
static void fdct32(const T *in, T *out) {
  T t[32];
  fdct_ii_32(in, t);
  for (int i = 0; i < 32; i++) out[i] = t[PERM32[i]];
}

/* forward.rs:1212-1277.  h[i] = halves at even i, A[i] = pairs at odd i. */
static void fdct_ii_32_asym(const T *h, const P *A, T *out) {
  T t[32], r[16];
  for (int i = 0; i < 16; i += 2) {
    bf_neg_asym(h[i], A[31 - i], &t[i], &t[31 - i]);
    bf_sub_asym(A[i + 1], h[30 - i], &t[i + 1], &t[30 - i]);
  }
  fdct_ii_16(t, out);
  for (int i = 0; i < 16; i++) r[i] = t[31 - i]; /* (tv, tu, tt, ..., tg) */
  fdst_iv_16(r, out + 16);
  rev(out + 16, 16);
}

/* stage-0 constants of daala_fdst_iv_32_asym, forward.rs:1286-1366, in call order */
static const int32_t ST0_32[16][6] = {
    {5933, 13, 22595, 14, 1137, 15},   {6203, 13, 21403, 14, 3409, 15},
    {25833, 15, 315, 8, 5673, 15},     {26791, 15, 4717, 12, 7923, 15},
    {6921, 13, 17531, 14, 10153, 15},  {28511, 15, 32303, 15, 1545, 12},
    {29269, 15, 14733, 14, 1817, 12},  {29957, 15, 13279, 14, 8339, 14},
    {7643, 13, 11793, 14, 18779, 15},  {15557, 14, 20557, 15, 20835, 15},
    {31581, 15, 17479, 15, 22841, 15}, {7993, 13, 14359, 15, 3099, 12},
    {16143, 14, 2801, 13, 26683, 15},  {16261, 14, 4011, 14, 14255, 14},
    {32679, 15, 4821, 15, 30269, 15},  {16379, 14, 201, 12, 15977, 14},
};

/* forward.rs:1279-1548.  A[i] = pairs at even i, h[i] = halves at odd i
 * (t0, t1h, t2, t3h, ..., tu, tvh). */
static void fdst_iv_32_asym(const P *A, const T *h, T *out) {
  T x[32];
  /* Stage 0: even i: (t_i, t_(31-i)) = RotateAdd::half(t_i, t_(31-i)h);
   *          odd  i: (t_(31-i), t_i) = RotateSub::half(t_(31-i), t_i h) */
  for (int i = 0; i < 16; i++) {
    const int32_t *m = ST0_32[i];
    if ((i & 1) == 0) rot_half(K_ADD, A[i], h[31 - i], m[0], m[1], m[2], m[3], m[4], m[5], &x[i], &x[31 - i]);
    else rot_half(K_SUB, A[31 - i], h[i], m[0], m[1], m[2], m[3], m[4], m[5], &x[31 - i], &x[i]);
  }
  T t0 = x[0], t1 = x[1], t2 = x[2], t3 = x[3], t4 = x[4], t5 = x[5], t6 = x[6], t7 = x[7];
  T t8 = x[8], t9 = x[9], ta = x[10], tb = x[11], tc = x[12], td = x[13], te = x[14], tf = x[15];
  T tg = x[16], th = x[17], ti = x[18], tj = x[19], tk = x[20], tl = x[21], tm = x[22], tn = x[23];
  T to = x[24], tp = x[25], tq = x[26], tr = x[27], ts = x[28], tt = x[29], tu = x[30], tv = x[31];
  T t0h, t1h, t2h, t3h, t4h, t5h, t6h, t7h, t8h, t9h, tah, tbh, tch, tdh, teh, tfh;
  T tgh, thh, tih, tjh, tkh, tlh, tmh, tnh, toh, tph, tqh, trh, tsh, tth, tuh, tvh;
  P T0, TV, TH, TE, T2, TT, TJ, TC, T4, TR, TL, TA, T6, TP, TN, T8;
  /* Stage 1, forward.rs:1369-1384 */
  bf_add(t0, tf, &T0, &tfh);
  bf_sub(tv, tg, &TV, &tgh);
  bf_add(th, tu, &TH, &tuh);
  bf_sub(te, t1, &TE, &t1h);
  bf_add(t2, td, &T2, &tdh);
  bf_sub(tt, ti, &TT, &tih);
  bf_add(tj, ts, &TJ, &tsh);
  bf_sub(tc, t3, &TC, &t3h);
  bf_add(t4, tb, &T4, &tbh);
  bf_sub(tr, tk, &TR, &tkh);
  bf_add(tl, tq, &TL, &tqh);
  bf_sub(ta, t5, &TA, &t5h);
  bf_add(t6, t9, &T6, &t9h);
  bf_sub(tp, tm, &TP, &tmh);
  bf_add(tn, to, &TN, &toh);
  bf_sub(t8, t7, &T8, &t7h);
  /* Stage 2, forward.rs:1387-1402 */
  bf_sub_asym(T0, t7h, &t0, &t7);
  bf_add_asym(TV, toh, &tv, &to);
  bf_sub_asym(TP, tuh, &tp, &tu);
  bf_add_asym(T6, t1h, &t6, &t1);
  bf_sub_asym(T2, t5h, &t2, &t5);
  bf_add_asym(TT, tqh, &tt, &tq);
  bf_sub_asym(TR, tsh, &tr, &ts);
  bf_add_asym(T4, t3h, &t4, &t3);
  bf_add_asym(T8, tgh, &t8, &tg);
  bf_sub_asym(TE, tmh, &te, &tm);
  bf_add_asym(TN, tfh, &tn, &tf);
  bf_sub_asym(TH, t9h, &th, &t9);
  bf_add_asym(TA, tih, &ta, &ti);
  bf_sub_asym(TC, tkh, &tc, &tk);
  bf_add_asym(TL, tdh, &tl, &td);
  bf_sub_asym(TJ, tbh, &tj, &tb);
  /* Stage 3, forward.rs:1408-1444 */
  rot(K_SUB, tf, tg, 17911, 14, 14699, 14, 803, 13, &tf, &tg);
  rot(K_ADD, th, te, 10217, 13, 5461, 13, 1189, 12, &th, &te);
  rot(K_ADD, ti, td, 5543, 12, 3363, 13, 7723, 14, &ti, &td);
  rot(K_SUB, tc, tj, 11529, 13, 2271, 14, 5197, 13, &tc, &tj);
  rot_neg(0, tb, tk, 11529, 13, 2271, 14, 5197, 13, &tb, &tk);
  rot_neg(0, ta, tl, 5543, 12, 3363, 13, 7723, 14, &ta, &tl);
  rot_neg(0, t9, tm, 10217, 13, 5461, 13, 1189, 12, &t9, &tm);
  rot_neg(0, t8, tn, 17911, 14, 14699, 14, 803, 13, &t8, &tn);
  /* Stage 4, forward.rs:1447-1462 */
  P T3, TS, TU, T1, TO, TQ, T7, T5, TB, TK, TM, T9, TF, TG, TI, TD;
  bf_sub(t3, t0, &T3, &t0h);
  bf_add(ts, tv, &TS, &tvh);
  bf_sub(tu, tt, &TU, &tth);
  bf_add(t1, t2, &T1, &t2h);
  bf_add(to, t4, &TO, &t4h);
  bf_sub(tq, t6, &TQ, &t6h);
  bf_add(t7, tr, &T7, &trh);
  bf_sub(t5, tp, &T5, &tph);
  to = TO.f; tq = TQ.f; t7 = T7.f; t5 = T5.f; /* ((_h, full), xh) keeps the full value */
  bf_sub(tb, t8, &TB, &t8h);
  bf_add(tk, tn, &TK, &tnh);
  bf_sub(tm, tl, &TM, &tlh);
  bf_add(t9, ta, &T9, &tah);
  bf_sub(tf, tc, &TF, &tch);
  bf_add(tg, tj, &TG, &tjh);
  bf_sub(ti, th, &TI, &thh);
  bf_add(td, te, &TD, &teh);
  /* Stage 5, forward.rs:1468-1483 */
  rot(K_ADD, to, t7, 301, 8, 1609, 11, 6393, 15, &to, &t7);
  rot(K_ADD, tph, t6h, 11363, 13, 9041, 15, 4551, 13, &tph, &t6h);
  rot_neg(0, t5, tq, 5681, 12, 9041, 15, 4551, 13, &t5, &tq);
  rot_neg(0, t4h, trh, 9633, 13, 12873, 14, 6393, 15, &t4h, &trh);
  /* Stage 6, forward.rs:1486-1501 */
  bf_add_asym(T1, t0h, &t1, &t0);
  bf_sub_asym(TU, tvh, &tu, &tv);
  bf_sub_asym(TS, t2h, &ts, &t2);
  bf_sub_asym(T3, tth, &t3, &tt);
  bf_add_asym(hp(t5), t4h, &t5, &t4);
  bf_sub_asym(hp(tq), trh, &tq, &tr);
  bf_add_asym(hp(t7), t6h, &t7, &t6);
  bf_sub_asym(hp(to), tph, &to, &tp);
  bf_add_asym(T9, t8h, &t9, &t8);
  bf_sub_asym(TM, tnh, &tm, &tn);
  bf_sub_asym(TK, tah, &tk, &ta);
  bf_sub_asym(TB, tlh, &tb, &tl);
  bf_add_asym(TI, tch, &ti, &tc);
  bf_add_asym(TD, tjh, &td, &tj);
  bf_add_asym(TF, teh, &tf, &te);
  bf_sub_asym(TG, thh, &tg, &th);
  /* Stage 7, forward.rs:1507-1542 */
  rot_neg(0, t2, tt, 669, 9, 8867, 14, 3135, 13, &t2, &tt);
  rot(K_ADD, ts, t3, 669, 9, 8867, 14, 3135, 13, &ts, &t3);
  rot_neg(0, ta, tl, 669, 9, 8867, 14, 3135, 13, &ta, &tl);
  rot(K_ADD, tk, tb, 669, 9, 8867, 14, 3135, 13, &tk, &tb);
  rot(K_ADD, tc, tj, 669, 9, 8867, 14, 3135, 13, &tc, &tj);
  rot_neg(0, ti, td, 669, 9, 8867, 14, 3135, 13, &ti, &td);
  rot_pi4(PI4_ADD, tu, t1, 5793, 12, 5793, 13, &tu, &t1);
  rot_pi4(PI4_ADD, tq, t5, 5793, 12, 5793, 13, &tq, &t5);
  rot_pi4(PI4_SUB, tp, t6, 5793, 12, 5793, 13, &tp, &t6);
  rot_pi4(PI4_ADD, tm, t9, 5793, 12, 5793, 13, &tm, &t9);
  rot_pi4(PI4_ADD, te, th, 5793, 12, 5793, 13, &te, &th);
  T o[32] = {t0, t1, t2, t3, t4, t5, t6, t7, t8, t9, ta, tb, tc, td, te, tf,
             tg, th, ti, tj, tk, tl, tm, tn, to, tp, tq, tr, ts, tt, tu, tv};
  memcpy(out, o, sizeof o);
  (void)t3h; (void)tbh; (void)tdh; (void)tkh; (void)tqh; (void)tsh; (void)tuh; (void)tfh;
}

/* ---- 64-point --------------------------------------------------------------- */
/* forward.rs:1551-1684 */
static void fdct64(const T *in, T *out) {
  P asym[32]; T half[32];
  for (int i = 0; i < 16; i++) { /* butterfly_pair(i), forward.rs:1559-1567 */
    int j = 2 * i;
    T ah, dh; P b, c;
    bf_neg(in[j], in[63 - j], &ah, &c);
    bf_add(in[j + 1], in[63 - j - 1], &b, &dh);
    half[i] = ah;
    half[31 - i] = dh;
    asym[i] = b;
    asym[31 - i] = c;
  }
  T tmp[64];
  T h32[32]; P A32[32];
  /* (half[0], asym[0], half[1], asym[1], ..., half[15], asym[15]) */
  for (int k = 0; k < 16; k++) { h32[2 * k] = half[k]; A32[2 * k + 1] = asym[k]; }
  fdct_ii_32_asym(h32, A32, tmp);
  /* (asym[31], half[31], asym[30], half[30], ..., asym[16], half[16]) */
  for (int k = 0; k < 16; k++) { A32[2 * k] = asym[31 - k]; h32[2 * k + 1] = half[31 - k]; }
  fdst_iv_32_asym(A32, h32, tmp + 32);
  rev(tmp + 32, 32);
  /* reorder_4(i, j), forward.rs:1661-1683 : j = 4-bit reversal of i */
  static const uint8_t BR4[16] = {0, 8, 4, 12, 2, 10, 6, 14, 1, 9, 5, 13, 3, 11, 7, 15};
  for (int i = 0; i < 16; i++) {
    int j = BR4[i];
    out[4 * i + 0] = tmp[j];
    out[4 * i + 1] = tmp[32 + j];
    out[4 * i + 2] = tmp[16 + j];
    out[4 * i + 3] = tmp[48 + j];
  }
}

/* ---- 1-D dispatch, forward.rs:1745-1768 ----------------------------------- */
enum { T1_DCT = 0, T1_ADST = 1, T1_ID = 2 };

int r1o_fwd_txfm1d(int type1d, int n, const int32_t *in, int32_t *out) {
  if (type1d == T1_DCT) {
    switch (n) {
      case 4: fdct4(in, out); return 0;
      case 8: fdct8(in, out); return 0;
      case 16: fdct16(in, out); return 0;
      case 32: fdct32(in, out); return 0;
      case 64: fdct64(in, out); return 0;
    }
  } else if (type1d == T1_ADST) {
    switch (n) {
      case 4: fdst_vii_4(in, out); return 0;
      case 8: fdst8(in, out); return 0;
      case 16: fdst16(in, out); return 0;
    }
  } else if (type1d == T1_ID) {
    if (n == 4 || n == 8 || n == 16 || n == 32) { /* forward.rs:1686-1700: plain copy */
      memmove(out, in, (size_t)n * sizeof(T));
      return 0;
    }
  }
  return -1; /* `unimplemented!` in the reference */
}

/* ---- 2-D driver -------------------------------------------------------------- */
static const uint8_t TXW[R1O_TX_SIZES_ALL] = {4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64};
static const uint8_t TXH[R1O_TX_SIZES_ALL] = {4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16};
int r1o_tx_width(int s) { return (s >= 0 && s < R1O_TX_SIZES_ALL) ? TXW[s] : -1; }
int r1o_tx_height(int s) { return (s >= 0 && s < R1O_TX_SIZES_ALL) ? TXH[s] : -1; }

/* forward.rs:22-40, rows indexed by shift_idx = (bd-8)/2 */
static const int8_t SH_A[3][3] = {{3, 0, 0}, {2, 0, 1}, {0, 0, 3}};
static const int8_t SH_B[3][3] = {{4, -1, 0}, {2, 0, 1}, {0, 0, 3}};
static const int8_t SH_C[3][3] = {{4, -2, 0}, {2, 0, 0}, {0, 0, 2}};
static const int8_t SH_D[3][3] = {{4, -1, -2}, {2, 0, -1}, {0, 0, 1}};
static const int8_t (*shift_table(int tx_size))[3] {
  switch (tx_size) {
    case R1O_TX_4X4: return SH_A;
    case R1O_TX_32X32: case R1O_TX_16X32: case R1O_TX_32X16: case R1O_TX_16X64: case R1O_TX_64X16:
      return SH_C;
    case R1O_TX_64X64: case R1O_TX_32X64: case R1O_TX_64X32: return SH_D;
    default: return SH_B;
  }
}

/* round_shift_array(arr, n, -entry), src/transform/mod.rs:499-521: a table
 * entry > 0 is a (wrapping) left shift, < 0 a rounding right shift. */
static inline T apply_shift(T v, int entry) {
  if (entry > 0) return (T)((uint32_t)v << entry);
  if (entry < 0) return (v + ((1 << -entry) >> 1)) >> -entry;
  return v;
}

/* column / row 1-D type and flip of a 2-D TxType, src/transform/mod.rs:196-213 */
static void tx_type_1d(int tx_type, int *col, int *col_flip, int *row, int *row_flip) {
  /* 0 DCT, 1 ADST, 2 FLIPADST, 3 ID */
  static const uint8_t C[R1O_TX_TYPES] = {0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3};
  static const uint8_t R[R1O_TX_TYPES] = {0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2};
  static const uint8_t K[4] = {T1_DCT, T1_ADST, T1_ADST, T1_ID};
  *col = K[C[tx_type]]; *col_flip = C[tx_type] == 2;
  *row = K[R[tx_type]]; *row_flip = R[tx_type] == 2;
}

static int valid_1d(int t, int n) {
  if (t == T1_DCT) return 1;
  if (t == T1_ADST) return n <= 16;
  return n <= 32;
}

int r1o_fht_valid(int tx_size, int tx_type) {
  if (tx_size < 0 || tx_size >= R1O_TX_SIZES_ALL || tx_type < 0 || tx_type >= R1O_TX_TYPES) return 0;
  int ct, cf, rt, rf;
  tx_type_1d(tx_type, &ct, &cf, &rt, &rf);
  return valid_1d(ct, TXH[tx_size]) && valid_1d(rt, TXW[tx_size]);
}

/* forward.rs:1804-1899, one column / one row at a time (the SIMD lane width of
 * the reference does not change any value). */
int r1o_fht(const int16_t *residual, int32_t *coeffs, int tx_size, int tx_type, int bd) {
  if (!r1o_fht_valid(tx_size, tx_type) || (bd != 8 && bd != 10 && bd != 12)) return -1;
  const int W = TXW[tx_size], H = TXH[tx_size];
  const int8_t *sh = shift_table(tx_size)[(bd - 8) / 2];
  int ct, cf, rt, rf;
  tx_type_1d(tx_type, &ct, &cf, &rt, &rf);
  static _Thread_local T buf[64 * 64];
  T a[64], b[64];
  for (int c = 0; c < W; c++) {
    for (int r = 0; r < H; r++) {
      int rr = cf ? (H - r - 1) : r; /* :1829-1834 flip upside down */
      a[r] = apply_shift((T)residual[rr * W + c], sh[0]);
    }
    r1o_fwd_txfm1d(ct, H, a, b);
    int cc = rf ? (W - c - 1) : c; /* intended semantic of :1858-1862 */
    for (int r = 0; r < H; r++) buf[r * W + cc] = apply_shift(b[r], sh[1]);
  }
  for (int r = 0; r < H; r++) {
    r1o_fwd_txfm1d(rt, W, buf + r * W, b);
    for (int c = 0; c < W; c++) coeffs[r * W + c] = apply_shift(b[c], sh[2]);
  }
  return 0;
}

/* src/encoder.rs:1045-1059 */
void r1o_diff(int16_t *dst, const void *s1, ptrdiff_t st1, const void *s2, ptrdiff_t st2, int w,
              int h, int pb) {
  for (int r = 0; r < h; r++)
    for (int c = 0; c < w; c++) {
      int32_t a = pb == 1 ? ((const uint8_t *)s1)[r * st1 + c] : ((const uint16_t *)s1)[r * st1 + c];
      int32_t b = pb == 1 ? ((const uint8_t *)s2)[r * st2 + c] : ((const uint16_t *)s2)[r * st2 + c];
      dst[r * w + c] = (int16_t)(a - b);
    }
}
