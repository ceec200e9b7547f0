// gba_solver.cuh -- reduced camera system: tile structure, fp64 tile Cholesky, solve.
//
// The reference factorises the 6P x 6P system on the HOST: D2H copy, fp32->fp64, triplet
// assembly and Eigen::SimplicialLLT<double> every iteration (droid_kernels.cu:1131-1213).
// Here the matrix lives on the device as fp64 NB x NB tiles (lower triangle, only tiles that
// are structurally non-zero after symbolic fill are stored / touched); the numeric result is
// that of a dense Cholesky, the work is that of the block-sparse one.
#pragma once
#include "gba_common.cuh"

#define TS_MAXT 512  // max tile rows (6 * P <= 32768)

// Mark the tiles touched by the pose pairs of one depth keyframe (its own pose and the
// destination poses of its out-edges are all mutually coupled through the Schur complement;
// the pose-pose Hessian blocks (ii,ii),(ii,jj),(jj,jj) are a subset).
__global__ void k_tile_mark(BaParams p, unsigned char *tmask) {
  const int k = blockIdx.x;
  if (k >= p.st->K) return;
  const int e0 = p.eptr[k], deg = p.eptr[k + 1] - e0;
  const int L = deg + 1;
  auto pose_of = [&](int q) {
    const int pr = (q < deg) ? (int)p.jj[p.eidx[e0 + q]] - p.t0 : p.kx[k] - p.t0;
    return (pr >= 0 && pr < p.P) ? pr : -1;
  };
  for (int idx = threadIdx.x; idx < L * L; idx += blockDim.x) {
    const int a = pose_of(idx / L), b = pose_of(idx % L);
    if (a < 0 || b < 0 || a < b) continue;
    const int r0 = (6 * a) / GBA_NB, r1 = (6 * a + 5) / GBA_NB;
    const int c0 = (6 * b) / GBA_NB, c1 = (6 * b + 5) / GBA_NB;
    for (int r = r0; r <= r1; r++)
      for (int c = c0; c <= c1; c++)
        if (r >= c) tmask[r * p.nT + c] = 1;
  }
}

// One CTA: symbolic fill of the tile structure (right-looking elimination on booleans), slot
// numbering, per-column lists of active sub-diagonal tiles.
__global__ void __launch_bounds__(1024) k_tile_fill(BaParams p, unsigned char *tmask, int *colrows, int *colcnt) {
  const int nT = p.nT, tid = threadIdx.x;
  __shared__ int s_cnt, s_max;
  for (int i = tid; i < nT; i += blockDim.x) tmask[i * nT + i] = 1;
  __syncthreads();
  for (int kk = 0; kk < nT; kk++) {
    const int m = nT - kk - 1;
    for (int idx = tid; idx < m * m; idx += blockDim.x) {
      const int i = kk + 1 + idx / m, j = kk + 1 + idx % m;
      if (j <= i && tmask[i * nT + kk] && tmask[j * nT + kk]) tmask[i * nT + j] = 1;
    }
    __syncthreads();
  }
  // slots in row-major lower order (single thread: nT^2/2 <= 131k steps, once per call)
  if (tid == 0) {
    int s = 0;
    for (int i = 0; i < nT; i++)
      for (int j = 0; j < nT; j++) p.tslot[i * nT + j] = (j <= i && tmask[i * nT + j]) ? s++ : -1;
    s_cnt = s;
    s_max = 0;
  }
  __syncthreads();
  for (int kk = tid; kk < nT; kk += blockDim.x) {
    int c = 0;
    for (int i = kk + 1; i < nT; i++)
      if (tmask[i * nT + kk]) colrows[kk * nT + c++] = i;
    colcnt[kk] = c;
    atomicMax(&s_max, c);
  }
  __syncthreads();
  if (tid == 0) {
    p.st->n_slots = s_cnt;
    p.st->max_col = s_max;
  }
}

// padded rows/cols of the last tile get a unit diagonal so the factorisation runs through
__global__ void k_pad_diag(BaParams p) {
  const int i = 6 * p.P + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < p.ld) *sys_at(p, i, i) = 1.0;
}

// --------------------------------------------------------------------------------------------
// Factorisation of one tile column kk: (1) Cholesky of the diagonal tile + its inverse,
// (2) L[i,kk] = A[i,kk] * inv(L_kk)^T for the active tiles below, (3) trailing update of the
// active tile pairs.  All fp64 (the reference solves in fp64, droid_kernels.cu:1133,1199).
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_potrf_diag(BaParams p, int kk, double *invL, int *fail) {
  extern __shared__ double s_buf[];
  double(*A)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf);
  double(*X)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf + GBA_NB * (GBA_NB + 1));
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  double *T = p.sys + (size_t)p.tslot[kk * p.nT + kk] * GBA_NB2;
  if (*fail) return;  // an earlier tile column already failed
  if (tid == 0) s_bad = 0;
  for (int i = tid; i < GBA_NB2; i += 256) {
    const int r = i / GBA_NB, c = i % GBA_NB;
    A[r][c] = (c <= r) ? T[i] : 0.0;
    X[r][c] = 0.0;
  }
  __syncthreads();
  for (int j = 0; j < GBA_NB; j++) {
    const double d = A[j][j];
    if (!(d > 0.0) || isinf(d)) {  // SimplicialLLT reports NumericalIssue on a non-positive pivot
      if (tid == 0) s_bad = 1;
      break;  // uniform: every thread reads the same d
    }
    const double inv = 1.0 / sqrt(d);
    __syncthreads();
    for (int i = j + tid; i < GBA_NB; i += 256) A[i][j] = (i == j) ? sqrt(d) : A[i][j] * inv;
    __syncthreads();
    // trailing update of the lower triangle
    const int m = GBA_NB - j - 1;
    for (int idx = tid; idx < m * m; idx += 256) {
      const int r = j + 1 + idx / m, c = j + 1 + idx % m;
      if (c <= r) A[r][c] -= A[r][j] * A[c][j];
    }
    __syncthreads();
  }
  __syncthreads();
  if (s_bad) {
    if (tid == 0) *fail = 1;
    return;
  }
  // X = inv(L): row i of X from the rows above, X[i][c] = -(sum_{m=c}^{i-1} L[i][m] X[m][c]) / L[i][i]
  for (int i = 0; i < GBA_NB; i++) {
    if (tid < i) {
      double s = 0.0;
      for (int m = tid; m < i; m++) s += A[i][m] * X[m][tid];
      X[i][tid] = -s / A[i][i];
    } else if (tid == i)
      X[i][i] = 1.0 / A[i][i];
    __syncthreads();
  }
  double *XI = invL + (size_t)kk * GBA_NB2;
  for (int i = tid; i < GBA_NB2; i += 256) {
    const int r = i / GBA_NB, c = i % GBA_NB;
    T[i] = (c <= r) ? A[r][c] : 0.0;
    XI[i] = X[r][c];
  }
}

// L[i,kk] = A[i,kk] * inv(L_kk)^T ; one CTA per active tile of column kk
__global__ void __launch_bounds__(256) k_trsm(BaParams p, int kk, const double *invL, const int *colrows,
                                              const int *colcnt, const int *fail) {
  extern __shared__ double s_buf[];
  double(*A)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf);
  double(*X)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf + GBA_NB * (GBA_NB + 1));
  if (*fail) return;
  if ((int)blockIdx.x >= colcnt[kk]) return;
  const int i = colrows[kk * p.nT + blockIdx.x], tid = threadIdx.x;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2;
  const double *XI = invL + (size_t)kk * GBA_NB2;
  for (int e = tid; e < GBA_NB2; e += 256) {
    A[e / GBA_NB][e % GBA_NB] = T[e];
    X[e / GBA_NB][e % GBA_NB] = XI[e];
  }
  __syncthreads();
  // out[r][c] = sum_m A[r][m] * X[c][m], m <= c (X lower triangular); 4x4 outputs per thread
  const int r0 = (tid / 16) * 4, c0 = (tid % 16) * 4;
  double acc[4][4] = {};
  for (int m = 0; m < GBA_NB; m++) {
    double a[4], x[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      a[u] = A[r0 + u][m];
      x[u] = X[c0 + u][m];
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int v = 0; v < 4; v++) acc[u][v] += a[u] * x[v];
  }
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) T[(r0 + u) * GBA_NB + c0 + v] = acc[u][v];
}

// A[i,j] -= L[i,kk] * L[j,kk]^T for active i >= j of column kk
__global__ void __launch_bounds__(256) k_syrk(BaParams p, int kk, const int *colrows, const int *colcnt,
                                              const int *fail) {
  extern __shared__ double s_buf[];
  double(*A)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf);
  double(*B)[GBA_NB + 1] = reinterpret_cast<double(*)[GBA_NB + 1]>(s_buf + GBA_NB * (GBA_NB + 1));
  if (*fail) return;
  const int cnt = colcnt[kk];
  // blockIdx.x -> (a >= b)
  int a = 0, rem = blockIdx.x;
  while (rem > a) {
    rem -= a + 1;
    a++;
  }
  const int b = rem;
  if (a >= cnt) return;
  const int i = colrows[kk * p.nT + a], j = colrows[kk * p.nT + b], tid = threadIdx.x;
  const double *Li = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2;
  const double *Lj = p.sys + (size_t)p.tslot[j * p.nT + kk] * GBA_NB2;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + j] * GBA_NB2;
  for (int e = tid; e < GBA_NB2; e += 256) {
    A[e / GBA_NB][e % GBA_NB] = Li[e];
    B[e / GBA_NB][e % GBA_NB] = Lj[e];
  }
  __syncthreads();
  const int r0 = (tid / 16) * 4, c0 = (tid % 16) * 4;
  double acc[4][4] = {};
  for (int m = 0; m < GBA_NB; m++) {
    double x[4], y[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      x[u] = A[r0 + u][m];
      y[u] = B[c0 + u][m];
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
      for (int v = 0; v < 4; v++) acc[u][v] += x[u] * y[v];
  }
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) T[(r0 + u) * GBA_NB + c0 + v] -= acc[u][v];
}

// --------------------------------------------------------------------------------------------
// Triangular solves L y = b, L^T x = y over the tile structure; one CTA, 256 threads
// (thread = (row, quarter)).  dx = (float) x, or zero when the factorisation failed
// (droid_kernels.cu:1202-1210, chol.py:13-15).
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_tri_solve(BaParams p, const double *invL, const int *colrows,
                                                   const int *colcnt, int *fail) {
  extern __shared__ double s_x[];  // [ld]
  __shared__ double s_y[GBA_NB];
  const int tid = threadIdx.x, row = tid >> 2, part = tid & 3, nT = p.nT;
  const double *rhs = p.sys + (size_t)p.st->n_slots * GBA_NB2;
  if (*fail) {
    for (int i = tid; i < 6 * p.P; i += 256) p.dx[i] = 0.f;
    if (tid == 0) {
      p.st->chol_fail += 1;
      *fail = 0;
    }
    return;
  }
  for (int i = tid; i < p.ld; i += 256) s_x[i] = rhs[i];
  __syncthreads();
  // forward
  for (int kk = 0; kk < nT; kk++) {
    const double *XI = invL + (size_t)kk * GBA_NB2 + row * GBA_NB + part * 16;
    double s = 0.0;
#pragma unroll
    for (int m = 0; m < 16; m++) s += XI[m] * s_x[kk * GBA_NB + part * 16 + m];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    __syncthreads();
    if (part == 0) {
      s_y[row] = s;
      s_x[kk * GBA_NB + row] = s;
    }
    __syncthreads();
    const int cnt = colcnt[kk];
    for (int a = 0; a < cnt; a++) {
      const int i = colrows[kk * nT + a];
      const double *L = p.sys + (size_t)p.tslot[i * nT + kk] * GBA_NB2 + row * GBA_NB + part * 16;
      double u = 0.0;
#pragma unroll
      for (int m = 0; m < 16; m++) u += L[m] * s_y[part * 16 + m];
      u += __shfl_xor_sync(0xffffffffu, u, 1);
      u += __shfl_xor_sync(0xffffffffu, u, 2);
      if (part == 0) s_x[i * GBA_NB + row] -= u;
    }
    __syncthreads();
  }
  // backward
  for (int kk = nT - 1; kk >= 0; kk--) {
    const int cnt = colcnt[kk];
    // t = y_kk - sum_i L[i,kk]^T x_i : thread (col = row, part) sums a quarter of the rows
    double t = 0.0;
    for (int a = 0; a < cnt; a++) {
      const int i = colrows[kk * nT + a];
      const double *L = p.sys + (size_t)p.tslot[i * nT + kk] * GBA_NB2;
#pragma unroll 4
      for (int m = part * 16; m < part * 16 + 16; m++) t += L[m * GBA_NB + row] * s_x[i * GBA_NB + m];
    }
    t += __shfl_xor_sync(0xffffffffu, t, 1);
    t += __shfl_xor_sync(0xffffffffu, t, 2);
    __syncthreads();
    if (part == 0) s_y[row] = s_x[kk * GBA_NB + row] - t;
    __syncthreads();
    // x_kk = inv(L_kk)^T * s_y
    const double *XI = invL + (size_t)kk * GBA_NB2;
    double s = 0.0;
#pragma unroll 4
    for (int m = part * 16; m < part * 16 + 16; m++) s += XI[m * GBA_NB + row] * s_y[m];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    __syncthreads();
    if (part == 0) s_x[kk * GBA_NB + row] = s;
    __syncthreads();
  }
  // the reference hands dx back as fp32 (droid_kernels.cu:1204-1205)
  for (int i = tid; i < 6 * p.P; i += 256) p.dx[i] = (float)s_x[i];
}
