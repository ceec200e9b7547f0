// gba_common.cuh -- shared types and per-pixel math of the B200 dense-BA kernels.
//
// Not a port of the reference's src/lib/droid_kernels.cu: the reference accumulates a
// 12x12 Hessian per pixel and materialises E rows in HBM (droid_kernels.cu:281-378); here
// only the 21+6 entries of (H_jj, v_j) are accumulated per edge (J_i = -Ad^T J_j is applied
// once per edge afterwards) and E rows never leave the SM.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GBA_NB 64            // tile size of the reduced camera matrix (fp64)
#define GBA_NB2 (GBA_NB * GBA_NB)

// Per-edge geometry, rebuilt from the current poses at the start of every iteration.
// Matches G_ij = T_j * T_i^-1 of droid_kernels.cu:96-107 / projective_ops.py:103.
struct __align__(16) EdgeGeo {
  float R[9];   // rotation of G_ij, row major
  float t[3];   // translation of G_ij
  int pi;       // ii - t0 if that pose is optimised, else -1
  int pj;       // jj - t0 if that pose is optimised, else -1
  int flags;    // bit0: stereo self edge (ii == jj): pose terms dropped (droid_kernels.cu:323,356)
  int pad;
};
static_assert(sizeof(EdgeGeo) == 64, "EdgeGeo must be 64 bytes");

struct Status {          // device -> host after the index build
  int K;                 // depth keyframes
  int max_deg;           // largest out-degree
  int err;               // bit0 index out of range, bit1 capacity, bit2 kx_global mismatch, bit3 a local edge
                         // needs a tile that the whole-graph edge list (ii_all / jj_all) does not have
  int n_tiles;           // tile rows
  int n_slots;           // active lower tiles after fill
  int max_col;           // max active sub-diagonal tiles in a tile column
  int chol_fail;         // running count of failed factorisations
  int max_row;           // max active tiles left of the diagonal in a tile row
  int n_closure;         // poses moved to the end by the arrowhead ordering
  int band_w;            // chosen long-range threshold (-1: none)
  int n_seg;             // independent band segments (nested dissection); 1 = plain band
  int seg_tiles;         // tile columns per segment
  int n_pos;             // pose positions of the reduced system incl. padding
  int cc_slot0;          // first slot of the closure x closure tiles (they are numbered last)
  int last_fail;         // code of the most recent failed factorisation: 1 pivot <= 0, 3 / 4 a pipeline watchdog
  int n_cc;              // keyframes whose out-degree exceeds tc_maxdeg: the CUDA-core Gram's work list (cc_list)
  int n_seg2;            // second dissection level: separators factorised side by side at the head of the closure block
  int seg2_tiles;        // tile columns of each of them
  unsigned long long gram_floats;  // total size of the per-frame Gram buffers
};

struct BaParams {
  // caller arrays
  float *poses, *disps;
  const float *intr, *sens, *targets, *weights, *eta;
  const int64_t *ii, *jj;
  int N, E, HW, wd, t0, t1, P, mode, motion_only, depth_only;
  float lm, ep;
  // index structures
  int *kx, *fmap, *deg, *eptr, *eidx, *ek;
  int *cc_list;  // [n_cc] keyframes of k_gram (k_prep_scan)
  const unsigned char *owned;
  unsigned long long *goff;
  Status *st;
  // per-iteration workspaces
  EdgeGeo *geo;
  float *hpart;   // [E][nchunk][32]
  float *edge_hv; // [E][32]
  int nchunk;
  float *Q, *W;   // [K][HW]
  float *gram;
  float *dx;      // [P][6]
  float *dz_out;
  // reduced system (tile packed, fp64)
  double *sys;    // [n_slots][NB*NB] then rhs [ld]
  int *tslot;     // [nT][nT]
  int *perm;      // [P] pose -> position in the reduced system (ordering below)
  int *iperm;     // [positions] position -> pose, -1 for padding positions
  int nT;         // STRIDE of the tile tables (upper bound); the actual count is st->n_tiles
  int *fail;      // the solver's failure flag (SolverBufs::fail): a pipeline watchdog raises it too
  int *wdog;      // [256] watchdog records of k_gram_tc: [0] = count, then 6 ints per record
  int il;         // targets / weights are [E,ht,wd,2] (interleaved, as the update network emits them) instead of [E,2,ht,wd]
  int tc_maxdeg;  // keyframes with out-degree <= this go to k_gram_tc (tensor cores), the rest to k_gram; 0 = all k_gram
  int ld;
};

__device__ __forceinline__ float ldg_stream(const float *p) { return __ldcs(p); }
__device__ __forceinline__ float4 ldg_stream4(const float *p) { return __ldcs(reinterpret_cast<const float4 *>(p)); }
// u / v components of 4 consecutive pixels px..px+3 of one edge's [2,HW] (planar) or [HW,2] (interleaved) block
__device__ __forceinline__ void ldg_uv4(const float *base, int HW, int px, bool il, float4 &u, float4 &v) {
  if (il) {
    const float4 a = ldg_stream4(base + 2 * px), b = ldg_stream4(base + 2 * px + 4);
    u = make_float4(a.x, a.z, b.x, b.z);
    v = make_float4(a.y, a.w, b.y, b.w);
  } else {
    u = ldg_stream4(base + px);
    v = ldg_stream4(base + HW + px);
  }
}

// address of element (r, c), r >= c, of the tile-packed lower-triangular system
__device__ __forceinline__ double *sys_at(const BaParams &p, int r, int c) {
  const int tr = r / GBA_NB, tc = c / GBA_NB;
  const int slot = p.tslot[tr * p.nT + tc];
  return p.sys + (size_t)slot * GBA_NB2 + (r % GBA_NB) * GBA_NB + (c % GBA_NB);
}

// Relative pose G_ij = T_j * T_i^-1 of two keyframes stored as (t, q xyzw): quaternion q and translation t
// (relSE3 + actSO3 of droid_kernels.cu:58-68,96-107; the only place this formula lives).
__device__ __forceinline__ void rel_se3(const float *Pi, const float *Pj, float (&q)[4], float (&t)[3]) {
  const float *qi = Pi + 3, *qj = Pj + 3;
  q[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
  q[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
  q[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
  q[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
  // t_ij = t_j - q_ij * t_i
  const float uv0 = 2.f * (q[1] * Pi[2] - q[2] * Pi[1]);
  const float uv1 = 2.f * (q[2] * Pi[0] - q[0] * Pi[2]);
  const float uv2 = 2.f * (q[0] * Pi[1] - q[1] * Pi[0]);
  t[0] = Pj[0] - (Pi[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1));
  t[1] = Pj[1] - (Pi[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2));
  t[2] = Pj[2] - (Pi[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0));
}
// rotation matrix (row major) equivalent to the quaternion sandwich X + w uv + q x uv, uv = 2 q x X
__device__ __forceinline__ void quat_to_rot(const float (&q)[4], float *R) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
  R[3] = 2.f * (x * y + w * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - w * x);
  R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
}

// y = B x with B = -Ad(G_ij)^T  (J_i = B J_j; droid_kernels.cu:79-94,325-326; se3.h:58-67,84-86)
template <typename T>
__device__ __forceinline__ void apply_B(const float *R, const float *t, const T *x, T *y) {
  // Ad^T [a;b] = [R^T a ; R^T (b - t x a)]
  const T a0 = x[0], a1 = x[1], a2 = x[2];
  const T c0 = x[3] - ((T)t[1] * a2 - (T)t[2] * a1);
  const T c1 = x[4] - ((T)t[2] * a0 - (T)t[0] * a2);
  const T c2 = x[5] - ((T)t[0] * a1 - (T)t[1] * a0);
  y[0] = -((T)R[0] * a0 + (T)R[3] * a1 + (T)R[6] * a2);
  y[1] = -((T)R[1] * a0 + (T)R[4] * a1 + (T)R[7] * a2);
  y[2] = -((T)R[2] * a0 + (T)R[5] * a1 + (T)R[8] * a2);
  y[3] = -((T)R[0] * c0 + (T)R[3] * c1 + (T)R[6] * c2);
  y[4] = -((T)R[1] * c0 + (T)R[4] * c1 + (T)R[7] * c2);
  y[5] = -((T)R[2] * c0 + (T)R[5] * c1 + (T)R[8] * c2);
}

// y = B^T x = -Ad(G_ij) x,  Ad [a;b] = [R a + t x (R b) ; R b]
__device__ __forceinline__ void apply_Bt(const float *R, const float *t, const float *x, float *y) {
  const float ra0 = R[0] * x[0] + R[1] * x[1] + R[2] * x[2];
  const float ra1 = R[3] * x[0] + R[4] * x[1] + R[5] * x[2];
  const float ra2 = R[6] * x[0] + R[7] * x[1] + R[8] * x[2];
  const float rb0 = R[0] * x[3] + R[1] * x[4] + R[2] * x[5];
  const float rb1 = R[3] * x[3] + R[4] * x[4] + R[5] * x[5];
  const float rb2 = R[6] * x[3] + R[7] * x[4] + R[8] * x[5];
  y[0] = -(ra0 + (t[1] * rb2 - t[2] * rb1));
  y[1] = -(ra1 + (t[2] * rb0 - t[0] * rb2));
  y[2] = -(ra2 + (t[0] * rb1 - t[1] * rb0));
  y[3] = -rb0;
  y[4] = -rb1;
  y[5] = -rb2;
}

// Linearisation of one edge pixel (Appendix A of SURVEY.md).
//   droid: droid_kernels.cu:290-319,345-352 (MIN_DEPTH 0.25: d = 0, w = 0 behind it)
//   geom : projective_ops.py:44,112 (valid = Z > 0.2; Z < 0.1 -> d = 1)
// J_j(u) = (ju0, 0, ju2, ju3, ju4, ju5), J_j(v) = (0, jv1, jv2, jv3, jv4, jv5).
struct PixLin {
  float ru, rv, wu, wv, jzu, jzv;
  float ju0, ju2, ju3, ju4, ju5;
  float jv1, jv2, jv3, jv4, jv5;
};

__device__ __forceinline__ void pix_linearize(const float *__restrict__ R, const float *__restrict__ t, float fx,
                                              float fy, float cx, float cy, int mode, float x0, float y0, float z,
                                              float tu, float tv, float wgu, float wgv, PixLin &o) {
  const float X = fmaf(R[0], x0, fmaf(R[1], y0, R[2])) + t[0] * z;
  const float Y = fmaf(R[3], x0, fmaf(R[4], y0, R[5])) + t[1] * z;
  const float Z = fmaf(R[6], x0, fmaf(R[7], y0, R[8])) + t[2] * z;
  bool invalid;
  float d;
  if (mode == 0) {
    invalid = Z < 0.25f;
    d = invalid ? 0.0f : __frcp_rn(Z);
  } else {
    invalid = !(Z > 0.2f);
    d = (Z < 0.1f) ? 1.0f : __frcp_rn(Z);
  }
  const float d2 = d * d;
  o.wu = invalid ? 0.0f : 0.001f * wgu;
  o.wv = invalid ? 0.0f : 0.001f * wgv;
  o.ru = tu - fmaf(fx * d, X, cx);
  o.rv = tv - fmaf(fy * d, Y, cy);
  const float fxd = fx * d, fyd = fy * d, fxd2 = fx * d2, fyd2 = fy * d2;
  o.ju0 = fxd * z;
  o.ju2 = -fxd2 * X * z;
  o.ju3 = -fxd2 * X * Y;
  o.ju4 = fmaf(fxd2 * X, X, fx);
  o.ju5 = -fxd * Y;
  o.jv1 = fyd * z;
  o.jv2 = -fyd2 * Y * z;
  o.jv3 = -fmaf(fyd2 * Y, Y, fy);
  o.jv4 = fyd2 * X * Y;
  o.jv5 = fyd * X;
  o.jzu = fxd * t[0] - fxd2 * X * t[2];
  o.jzv = fyd * t[1] - fyd2 * Y * t[2];
}

// ---------------------------------------------------------------------------------------------
// Lean form of the same linearisation for the hot kernels.  The focal lengths are folded into the
// weights (wu' = 0.001 fx^2 weight_u, wv' likewise) and the residuals (ru' = ru / fx), so that the
// Jacobian rows become  ju' = (e, 0, -a e, -a b, 1 + a^2, -b),  jv' = (0, e, -b e, -(1 + b^2), a b, a)
// with a = X/Z, b = Y/Z, e = z/Z, and  jz' = ((tx - tz a)/Z, (ty - tz b)/Z).  Every product of the
// normal equations keeps its value:  w J J^T = w' j' j'^T,  w r J = w' r' j',  w Jz J = w' jz' j'.
// ---------------------------------------------------------------------------------------------
struct PixConst {  // per launch
  float ifx, ify, cxf, cyf, kwu, kwv;
};
__device__ __forceinline__ PixConst make_pix_const(const float *intr) {
  PixConst k;
  k.ifx = 1.0f / intr[0];
  k.ify = 1.0f / intr[1];
  k.cxf = intr[2] * k.ifx;
  k.cyf = intr[3] * k.ify;
  k.kwu = 0.001f * intr[0] * intr[0];
  k.kwv = 0.001f * intr[1] * intr[1];
  return k;
}
struct PixHat {
  float ru, rv, wu, wv, jzu, jzv;  // residuals / fx, fy; folded weights; jz'
  float e, nae, nab, opa, nb;      // ju' = (e, 0, nae, nab, opa, nb)
  float nbe, nopb, ab, a;          // jv' = (0, e, nbe, nopb, ab, a)
};
template <int MODE>
__device__ __forceinline__ void pix_hat(const float *__restrict__ R, const float *__restrict__ t, const PixConst &k,
                                        float x0, float y0, float z, float tu, float tv, float wgu, float wgv,
                                        PixHat &o) {
  const float X = fmaf(R[0], x0, fmaf(R[1], y0, fmaf(t[0], z, R[2])));
  const float Y = fmaf(R[3], x0, fmaf(R[4], y0, fmaf(t[1], z, R[5])));
  const float Z = fmaf(R[6], x0, fmaf(R[7], y0, fmaf(t[2], z, R[8])));
  bool invalid;
  float d;
  if (MODE == 0) {  // droid_kernels.cu:26,302-306
    invalid = Z < 0.25f;
    d = invalid ? 0.0f : __fdividef(1.0f, Z);
  } else {  // projective_ops.py:44,112
    invalid = !(Z > 0.2f);
    d = (Z < 0.1f) ? 1.0f : __fdividef(1.0f, Z);
  }
  const float a = X * d, b = Y * d, e = z * d;
  o.wu = invalid ? 0.0f : k.kwu * wgu;
  o.wv = invalid ? 0.0f : k.kwv * wgv;
  o.ru = fmaf(tu, k.ifx, -k.cxf) - a;
  o.rv = fmaf(tv, k.ify, -k.cyf) - b;
  o.e = e;
  o.a = a;
  o.nb = -b;
  o.ab = a * b;
  o.nab = -o.ab;
  o.nae = -(a * e);
  o.nbe = -(b * e);
  o.opa = fmaf(a, a, 1.0f);
  o.nopb = -fmaf(b, b, 1.0f);
  o.jzu = d * fmaf(-t[2], a, t[0]);
  o.jzv = d * fmaf(-t[2], b, t[1]);
}

// packed fp32x2 arithmetic (Blackwell FFMA2 / FMUL2): two lanes of work in ONE issue slot
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void fma2(unsigned long long &d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma2r(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ unsigned long long dup2(float v) { return pack2(v, v); }

// ---------------------------------------------------------------------------------------------
// pix_hat for TWO pixels of one edge at once: every fp32 operation of the
// scalar version becomes one packed f32x2 instruction on the pair -- the same roundings in the same order, half
// the issue slots (these kernels are bound by instruction issue, not by HBM).  Signs are not materialised:
// ju' = (e, 0, -ae, -ab, opa, -b), jv' = (0, e, -be, -opb, ab, a); consumers negate the constant operand instead.
// ---------------------------------------------------------------------------------------------
struct EdgeConst2 {  // per edge, packed (v, v)
  unsigned long long R0, R1, R2, R3, R4, R5, R6, R7, R8, t0, t1, t2, nt2;
};
__device__ __forceinline__ EdgeConst2 make_edge_const2(const float *R, const float *t) {
  EdgeConst2 c;
  c.R0 = dup2(R[0]); c.R1 = dup2(R[1]); c.R2 = dup2(R[2]); c.R3 = dup2(R[3]); c.R4 = dup2(R[4]); c.R5 = dup2(R[5]);
  c.R6 = dup2(R[6]); c.R7 = dup2(R[7]); c.R8 = dup2(R[8]); c.t0 = dup2(t[0]); c.t1 = dup2(t[1]); c.t2 = dup2(t[2]);
  c.nt2 = dup2(-t[2]);
  return c;
}
struct PixHat2 {
  unsigned long long wu, wv, jzu, jzv;           // folded weights (0 where invalid), jz'
  unsigned long long e, a, b, ae, ab, be, opa, opb;
  unsigned long long X, Y, d;                     // for the residuals (pix_res2)
};
template <int MODE>
__device__ __forceinline__ void pix_hat2(const EdgeConst2 &c, const PixConst &k, unsigned long long x2,
                                         unsigned long long y2, unsigned long long z2, unsigned long long wgu2,
                                         unsigned long long wgv2, PixHat2 &o) {
  o.X = fma2r(c.R0, x2, fma2r(c.R1, y2, fma2r(c.t0, z2, c.R2)));
  o.Y = fma2r(c.R3, x2, fma2r(c.R4, y2, fma2r(c.t1, z2, c.R5)));
  const unsigned long long Z = fma2r(c.R6, x2, fma2r(c.R7, y2, fma2r(c.t2, z2, c.R8)));
  float Z0, Z1;
  unpack2(Z, Z0, Z1);
  bool i0, i1;
  float d0, d1;
  if (MODE == 0) {  // droid_kernels.cu:26,302-306
    i0 = Z0 < 0.25f; i1 = Z1 < 0.25f;
    d0 = i0 ? 0.0f : __fdividef(1.0f, Z0);
    d1 = i1 ? 0.0f : __fdividef(1.0f, Z1);
  } else {  // projective_ops.py:44,112
    i0 = !(Z0 > 0.2f); i1 = !(Z1 > 0.2f);
    d0 = (Z0 < 0.1f) ? 1.0f : __fdividef(1.0f, Z0);
    d1 = (Z1 < 0.1f) ? 1.0f : __fdividef(1.0f, Z1);
  }
  o.d = pack2(d0, d1);
  o.a = mul2(o.X, o.d);
  o.b = mul2(o.Y, o.d);
  o.e = mul2(z2, o.d);
  const unsigned long long gate = pack2(i0 ? 0.0f : 1.0f, i1 ? 0.0f : 1.0f);
  // invalid ? 0 : kw * wg  ==  (kw * wg) * gate exactly (gate is 0 or 1; kw * wg is finite)
  o.wu = mul2(mul2(dup2(k.kwu), wgu2), gate);
  o.wv = mul2(mul2(dup2(k.kwv), wgv2), gate);
  o.ab = mul2(o.a, o.b);
  o.ae = mul2(o.a, o.e);
  o.be = mul2(o.b, o.e);
  const unsigned long long one = dup2(1.0f);
  o.opa = fma2r(o.a, o.a, one);
  o.opb = fma2r(o.b, o.b, one);
  o.jzu = mul2(o.d, fma2r(c.nt2, o.a, c.t0));
  o.jzv = mul2(o.d, fma2r(c.nt2, o.b, c.t1));
}
// residuals / fx, fy of the pair: ru' = (tu * ifx - cx/fx) - a
__device__ __forceinline__ void pix_res2(const PixHat2 &o, const PixConst &k, unsigned long long tu2,
                                         unsigned long long tv2, unsigned long long &ru, unsigned long long &rv) {
  const unsigned long long m1 = dup2(-1.0f);
  ru = fma2r(o.a, m1, fma2r(tu2, dup2(k.ifx), dup2(-k.cxf)));
  rv = fma2r(o.b, m1, fma2r(tv2, dup2(k.ify), dup2(-k.cyf)));
}

// Sum NV (<= 32) per-lane values across the warp so that lane l ends up holding the total of
// v[l]: 31 shuffles instead of 5 per value (butterfly that halves the live values per stage).
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool upper = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; i++) {
      const float send = upper ? v[i] : v[i + s];
      const float keep = upper ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}
