// gba_potrf.cuh -- Cholesky of one 64 x 64 fp64 diagonal tile on ONE CTA (256 threads), latency first.
//
// The tile Cholesky of the reduced camera system (gba_solver.cuh) is a dependency chain of diagonal
// tiles; nothing else on the GPU can start until each of them is done, so what counts here is the
// length of the chain of dependent fp64 operations (~40 cycles each on B200), not throughput.
// Replaces the numeric core of Eigen::SimplicialLLT<double> on the host (droid_kernels.cu:1199-1203).
//
// Two-level: A = [A11 .; A21 A22] with 32 x 32 blocks.
//   1. warp 0 factorises A11 = Lh D Lh^T in REGISTERS: lane i owns row i of the (symmetric) block and row
//      i of M = Lh^-1 (Gauss-Jordan on [A | I]); rows travel by warp shuffle, there is no barrier and no
//      shared-memory round trip inside the pivot loop.  Two pivots per round with one reciprocal chain
//      (1/d1 = d0 / (d0 f - e^2) does not wait for 1/d0).  Output: M11, d[0..31];  X11 = L11^-1 = D^-1/2 M11.
//   2. all warps: W = A21 M11^T, A22' = A22 - W D^-1 W^T            (two 32^3 products)
//   3. warp 0 factorises A22' the same way -> M22, d[32..63]
//   4. all warps: X21 = -X22 (L21 X11) = -D2^-1/2 M22 W D1^-1 M11    (two 32^3 products)
// Result: X = L^-1 (what the triangular solves and the substitutions consume; L itself is never needed
// again) and the forward-substituted rhs block y = X b.
#pragma once
#include "gba_common.cuh"

#define PF_N 32
#define PF_LD 33  // pitch of the 32 x 32 shared-memory blocks (doubles)

__device__ __forceinline__ double pf_rcp(double d) {  // ~1e-12 relative: seed + one Newton step (see fast_rcp)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  const double e = fma(-d, r, 1.0);
  return fma(r, e, r);
}

// One warp.  a[k] = A[lane][k] (full symmetric row).  On return m[c] = M[lane][c] for c < lane (unit diagonal
// implied, zeros above), the pivots are in s_d[0..31]; returns false if a pivot is not positive and finite.
__device__ __forceinline__ bool warp_ldl32(double (&a)[PF_N], double (&m)[PF_N], double *s_d, int lane) {
  bool bad = false;
#pragma unroll
  for (int c = 0; c < PF_N; c++) m[c] = 0.0;
#pragma unroll
  for (int j = 0; j < PF_N; j += 2) {
    // rows j and j+1 as they stand after pivots < j (by symmetry: columns j, j+1); the three entries the
    // reciprocal chain needs come first, the others are fetched where they are consumed
    const double d0 = __shfl_sync(0xffffffffu, a[j], j), e = __shfl_sync(0xffffffffu, a[j + 1], j);
    const double f = __shfl_sync(0xffffffffu, a[j + 1], j + 1);
    const double t = fma(d0, f, -e * e);  // d0 * d1
    const double dinv0 = pf_rcp(d0), rt = pf_rcp(t);
    const double mm = e * dinv0;          // Lh[j+1][j]
    const double d1 = fma(-e, mm, f);
    const double dinv1 = d0 * rt;
    if (!(d0 > 0.0) || isinf(d0) || !(d1 > 0.0) || isinf(d1)) bad = true;
    if (lane == 0) {
      s_d[j] = d0;
      s_d[j + 1] = d1;
    }
    // this lane's multipliers (rows below the pivot pair only)
    const bool below = lane > j + 1;
    const double ra = below ? a[j] * dinv0 : 0.0;
    const double rb = below ? fma(-a[j], mm, a[j + 1]) * dinv1 : 0.0;
#pragma unroll
    for (int k = j + 2; k < PF_N; k++) {
      const double ak = __shfl_sync(0xffffffffu, a[k], j);
      const double bk = fma(-ak, mm, __shfl_sync(0xffffffffu, a[k], j + 1));  // row j+1 after pivot j
      a[k] = fma(-rb, bk, fma(-ra, ak, a[k]));
    }
    // the same row operations on M (initially I): rows j, j+1 of M
#pragma unroll
    for (int c = 0; c < j; c++) {
      const double mA = __shfl_sync(0xffffffffu, m[c], j);
      const double mB = fma(-mm, mA, __shfl_sync(0xffffffffu, m[c], j + 1));
      m[c] = (lane == j + 1) ? mB : fma(-rb, mB, fma(-ra, mA, m[c]));
    }
    // column j: M_j[j] = 1, M_{j+1}[j] = -mm;  column j+1: M_{j+1}[j+1] = 1
    m[j] = (lane == j + 1) ? -mm : fma(rb, mm, -ra);
    m[j + 1] = -rb;
  }
  return !__any_sync(0xffffffffu, bad);
}

// Shared memory: 6 blocks of 32 x 33 doubles + 64 pivots + 64 rhs  (~52 KB)
struct PotrfSmem {
  double A21[PF_N][PF_LD], A22[PF_N][PF_LD], M11[PF_N][PF_LD], M22[PF_N][PF_LD], W[PF_N][PF_LD], V[PF_N][PF_LD];
  double d[2 * PF_N], rd[2 * PF_N], b[2 * PF_N];
  int bad;
};

// T: the diagonal tile in global memory (row major, lower triangle valid), XI: out, inverse of its Cholesky
// factor (row major, full 64 x 64 with zeros above the diagonal), b: rhs block in/out (y = X b).
// Returns false (every thread) when the tile is not positive definite; nothing is written then.
#ifdef PF_TIMERS
#define PF_T(i) if (threadIdx.x == 0) pf_t[i] = clock64();
__device__ long long pf_t[16];
#else
#define PF_T(i)
#endif
__device__ bool potrf_tile64(const double *T, double *XI, double *b, PotrfSmem &s) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  PF_T(0)
  if (tid == 0) s.bad = 0;
  if (tid < 2 * PF_N) s.b[tid] = __ldcg(b + tid);
  double a[PF_N], m[PF_N];
  if (warp == 0) {
    // row `lane` of A11, mirrored across the diagonal (two coalescing-friendly passes over L2)
#pragma unroll
    for (int k = 0; k < PF_N; k++) a[k] = __ldcg(T + (k <= lane ? lane * GBA_NB + k : k * GBA_NB + lane));
  } else {
    // A21 (rows 32..63, cols 0..31) and the lower triangle of A22 into shared memory
    for (int i = tid - 32; i < PF_N * PF_N; i += 224) {
      const int r = i >> 5, c = i & 31;
      s.A21[r][c] = __ldcg(T + (PF_N + r) * GBA_NB + c);
      const double v = (c <= r) ? __ldcg(T + (PF_N + r) * GBA_NB + PF_N + c) : 0.0;
      if (c <= r) {
        s.A22[r][c] = v;
        s.A22[c][r] = v;
      }
    }
  }
  __syncthreads();
  PF_T(1)
  if (warp == 0) {
    const bool ok = warp_ldl32(a, m, s.d, lane);
    if (!ok && lane == 0) s.bad = 1;
#pragma unroll
    for (int c = 0; c < PF_N; c++) s.M11[lane][c] = (c < lane) ? m[c] : (c == lane ? 1.0 : 0.0);
  }
  __syncthreads();
  PF_T(2)
  if (s.bad) return false;
  // thread -> 4 outputs (row r, columns c0 + 8 q): a warp reads one row of the left operand (broadcast)
  // and 32 consecutive... (pitch 33: conflict free both ways)
  const int r = tid >> 3, c0 = tid & 7;
  {  // W = A21 M11^T  (M11 unit lower: sum over k <= c)
    double acc[4][2] = {};
#pragma unroll 4
    for (int k = 0; k < PF_N; k += 2) {
      const double x0 = s.A21[r][k], x1 = s.A21[r][k + 1];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        acc[q][0] = fma(x0, s.M11[c0 + 8 * q][k], acc[q][0]);
        acc[q][1] = fma(x1, s.M11[c0 + 8 * q][k + 1], acc[q][1]);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++) s.W[r][c0 + 8 * q] = acc[q][0] + acc[q][1];
  }
  if (tid < PF_N) s.rd[tid] = 1.0 / s.d[tid];
  __syncthreads();
  PF_T(3)
  {  // A22' = A22 - W D1^-1 W^T   (full symmetric block)
    double acc[4][2] = {};
#pragma unroll 4
    for (int k = 0; k < PF_N; k += 2) {
      const double x0 = s.W[r][k] * s.rd[k], x1 = s.W[r][k + 1] * s.rd[k + 1];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        acc[q][0] = fma(x0, s.W[c0 + 8 * q][k], acc[q][0]);
        acc[q][1] = fma(x1, s.W[c0 + 8 * q][k + 1], acc[q][1]);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++) s.A22[r][c0 + 8 * q] -= acc[q][0] + acc[q][1];
  }
  __syncthreads();
  PF_T(4)
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < PF_N; k++) a[k] = s.A22[lane][k];
    const bool ok = warp_ldl32(a, m, s.d + PF_N, lane);
    if (!ok && lane == 0) s.bad = 1;
#pragma unroll
    for (int c = 0; c < PF_N; c++) s.M22[lane][c] = (c < lane) ? m[c] : (c == lane ? 1.0 : 0.0);
  } else {
    // meanwhile: V = W D1^-1 M11   (V[r][c] = sum_k W[r][k] rd[k] M11[k][c]), 224 threads
    for (int i = tid - 32; i < PF_N * PF_N; i += 224) {
      const int rr = i >> 5, cc = i & 31;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll 4
      for (int k = 0; k < PF_N; k += 2) {
        s0 = fma(s.W[rr][k] * s.rd[k], s.M11[k][cc], s0);
        s1 = fma(s.W[rr][k + 1] * s.rd[k + 1], s.M11[k + 1][cc], s1);
      }
      s.V[rr][cc] = s0 + s1;
    }
  }
  __syncthreads();
  PF_T(5)
  if (s.bad) return false;
  if (tid < 2 * PF_N) s.rd[tid] = rsqrt(s.d[tid]);  // 1 / L[i][i]
  __syncthreads();
  // X = [D1^-1/2 M11, 0; -D2^-1/2 M22 V, D2^-1/2 M22]
  {
    double acc[4][2] = {};
#pragma unroll 4
    for (int k = 0; k < PF_N; k += 2) {
      const double x0 = s.M22[r][k], x1 = s.M22[r][k + 1];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        acc[q][0] = fma(x0, s.V[k][c0 + 8 * q], acc[q][0]);
        acc[q][1] = fma(x1, s.V[k + 1][c0 + 8 * q], acc[q][1]);
      }
    }
    const double sc1 = s.rd[r], sc2 = s.rd[PF_N + r];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int c = c0 + 8 * q;
      const double x11 = sc1 * s.M11[r][c], x22 = sc2 * s.M22[r][c], x21 = -sc2 * (acc[q][0] + acc[q][1]);
      XI[r * GBA_NB + c] = x11;
      XI[r * GBA_NB + PF_N + c] = 0.0;
      XI[(PF_N + r) * GBA_NB + c] = x21;
      XI[(PF_N + r) * GBA_NB + PF_N + c] = x22;
      // keep X in shared memory for the rhs: reuse W (X11), A21 (X21), A22 (X22)
      s.W[r][c] = x11;
      s.A21[r][c] = x21;
      s.A22[r][c] = x22;
    }
  }
  __syncthreads();
  PF_T(6)
  if (tid < 2 * PF_N) {  // y = X b
    const int i = tid;
    double s0 = 0.0, s1 = 0.0;
    if (i < PF_N) {
#pragma unroll 4
      for (int k = 0; k < PF_N; k += 2) {
        s0 = fma(s.W[i][k], s.b[k], s0);
        s1 = fma(s.W[i][k + 1], s.b[k + 1], s1);
      }
    } else {
#pragma unroll 4
      for (int k = 0; k < PF_N; k += 2) {
        s0 = fma(s.A21[i - PF_N][k], s.b[k], s0);
        s1 = fma(s.A21[i - PF_N][k + 1], s.b[k + 1], s1);
        s0 = fma(s.A22[i - PF_N][k], s.b[PF_N + k], s0);
        s1 = fma(s.A22[i - PF_N][k + 1], s.b[PF_N + k + 1], s1);
      }
    }
    b[i] = s0 + s1;
  }
  PF_T(7)
  return true;
}
