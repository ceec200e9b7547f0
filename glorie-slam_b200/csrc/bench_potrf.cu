// bench_potrf.cu -- development harness (not part of libgba.so): the 64 x 64 diagonal-tile Cholesky on one CTA,
// new (gba_potrf.cuh) against old (role_potrf in gba_solver.cuh): residual || X A X^T - I || and cycles per tile.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o bench_potrf bench_potrf.cu && ./bench_potrf
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#define PF_TIMERS
// #define DM_TIMERS
#include "gba_solver.cuh"
#include "gba_potrf.cuh"
#include "gba_dmma.cuh"

__global__ void __launch_bounds__(256) k_dm(const double *A, double *T, double *XI, double *b, int reps, long long *cyc, int *ok) {
  extern __shared__ double s_buf[];
  DmPotrfSmem &s = *reinterpret_cast<DmPotrfSmem *>(s_buf);
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
    for (int i = threadIdx.x; i < GBA_NB2; i += 256) T[i] = A[i];
    if (threadIdx.x < 64) b[threadIdx.x] = 1.0 + threadIdx.x;
    __threadfence();
    __syncthreads();
    const long long t0 = clock64();
    const bool good = dm_potrf64(T, XI, b, s);
    __syncthreads();
    tot += clock64() - t0;
    if (threadIdx.x == 0) *ok = good;
  }
  if (threadIdx.x == 0) *cyc = tot / reps;
}


__global__ void __launch_bounds__(256) k_new(const double *A, double *T, double *XI, double *b, int reps, long long *cyc, int *ok) {
  extern __shared__ double s_buf[];
  PotrfSmem &s = *reinterpret_cast<PotrfSmem *>(s_buf);
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
    for (int i = threadIdx.x; i < GBA_NB2; i += 256) T[i] = A[i];
    if (threadIdx.x < 64) b[threadIdx.x] = 1.0 + threadIdx.x;
    __threadfence();
    __syncthreads();
    const long long t0 = clock64();
    const bool good = potrf_tile64(T, XI, b, s);
    __syncthreads();
    tot += clock64() - t0;
    if (threadIdx.x == 0) *ok = good;
  }
  if (threadIdx.x == 0) *cyc = tot / reps;
}

__global__ void __launch_bounds__(256) k_old(BaParams p, SolverBufs sb, const double *A, int reps, long long *cyc) {
  extern __shared__ double s_buf[];
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
    for (int i = threadIdx.x; i < GBA_NB2; i += 256) p.sys[i] = A[i];
    if (threadIdx.x < 64) p.sys[GBA_NB2 + threadIdx.x] = 1.0 + threadIdx.x;
    __threadfence();
    __syncthreads();
    const long long t0 = clock64();
    role_potrf(p, sb, 0, s_buf);
    __syncthreads();
    tot += clock64() - t0;
  }
  if (threadIdx.x == 0) *cyc = tot / reps;
}

int main(int argc, char **argv) {
  const int n = 64, reps = 50;
  for (int trial = 0; trial < 3; trial++) {
    std::vector<double> B(n * n), A(n * n, 0.0);
    srand(1 + trial);
    const double cond_boost = trial == 0 ? 1.0 : (trial == 1 ? 1e-3 : 1e-6);
    for (auto &v : B) v = rand() / (double)RAND_MAX - 0.5;
    for (int i = 0; i < n; i++)
      for (int j = 0; j <= i; j++) {
        double s = 0;
        for (int k = 0; k < n; k++) s += B[i * n + k] * B[j * n + k];
        A[i * n + j] = s + (i == j ? cond_boost : 0.0);  // lower triangle only
      }
    double *dA, *dT, *dX, *db, *dsys, *dinvL;
    long long *dc;
    int *dok, *dfail, *dslot;
    Status *dst;
    cudaMalloc(&dA, n * n * 8); cudaMalloc(&dT, n * n * 8); cudaMalloc(&dX, n * n * 8); cudaMalloc(&db, 64 * 8);
    cudaMalloc(&dsys, (n * n + 64) * 8); cudaMalloc(&dinvL, n * n * 8); cudaMalloc(&dc, 8); cudaMalloc(&dok, 4);
    cudaMalloc(&dfail, 4); cudaMalloc(&dslot, 4); cudaMalloc(&dst, sizeof(Status));
    cudaMemset(dfail, 0, 4); cudaMemset(dslot, 0, 4);
    Status st = {}; st.n_slots = 1; st.n_tiles = 1;
    cudaMemcpy(dst, &st, sizeof(st), cudaMemcpyHostToDevice);
    cudaMemcpy(dA, A.data(), n * n * 8, cudaMemcpyHostToDevice);
    const int smem_new = sizeof(PotrfSmem), smem_old = (2 * GBA_NB + 16) * (GBA_NB + 1) * 8;
    cudaFuncSetAttribute(k_new, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_new);
    cudaFuncSetAttribute(k_old, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_old);
    k_new<<<1, 256, smem_new>>>(dA, dT, dX, db, reps, dc, dok);
    cudaError_t e = cudaDeviceSynchronize();
    { long long t[16]; cudaMemcpyFromSymbol(t, pf_t, sizeof(t)); printf("phases:"); for (int i = 1; i < 8; i++) printf(" %lld", t[i] - t[i-1]); printf("\n"); }
    long long cn = 0, co = 0; int ok = 0;
    cudaMemcpy(&cn, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&ok, dok, 4, cudaMemcpyDeviceToHost);
    {
      cudaFuncSetAttribute(k_dm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DmPotrfSmem));
      k_dm<<<1, 256, sizeof(DmPotrfSmem)>>>(dA, dT, dX, db, reps, dc, dok);
      cudaError_t e3 = cudaDeviceSynchronize();
      long long cd = 0; int okd = 0;
      cudaMemcpy(&cd, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&okd, dok, 4, cudaMemcpyDeviceToHost);
      printf("dmma potrf: %lld cycles ok=%d %s smem %d\n", cd, okd, cudaGetErrorString(e3), (int)sizeof(DmPotrfSmem));
    }
    std::vector<double> X(n * n), y(64), Xo(n * n), yo(64);
    cudaMemcpy(X.data(), dX, n * n * 8, cudaMemcpyDeviceToHost); cudaMemcpy(y.data(), db, 64 * 8, cudaMemcpyDeviceToHost);
    BaParams p = {}; p.sys = dsys; p.tslot = dslot; p.nT = 1; p.st = dst;
    SolverBufs sb = {}; sb.invL = dinvL; sb.fail = dfail;
    k_old<<<1, 256, smem_old>>>(p, sb, dA, reps, dc);
    cudaError_t e2 = cudaDeviceSynchronize();
    cudaMemcpy(&co, dc, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(Xo.data(), dinvL, n * n * 8, cudaMemcpyDeviceToHost); cudaMemcpy(yo.data(), dsys + n * n, 64 * 8, cudaMemcpyDeviceToHost);
    auto resid = [&](const std::vector<double> &Xm) {  // || X A X^T - I ||_max with the full symmetric A
      double worst = 0;
      std::vector<double> XA(n * n, 0.0);
      for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double s = 0; for (int k = 0; k < n; k++) s += Xm[i * n + k] * (k >= j ? A[k * n + j] : A[j * n + k]); XA[i * n + j] = s; }
      for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double s = 0; for (int k = 0; k < n; k++) s += XA[i * n + k] * Xm[j * n + k]; worst = fmax(worst, fabs(s - (i == j))); }
      return worst;
    };
    double dxy = 0, dX_ = 0, nX = 0;
    for (int i = 0; i < 64; i++) dxy = fmax(dxy, fabs(y[i] - yo[i]) / (fabs(yo[i]) + 1e-300));
    for (int i = 0; i < n * n; i++) { dX_ = fmax(dX_, fabs(X[i] - Xo[i])); nX = fmax(nX, fabs(Xo[i])); }
    printf("trial %d (diag boost %g): dmma(X shown) / warp-shuffle %lld cycles (ok=%d, %s), old %lld cycles (%s); resid new %.2e old %.2e; max|Xnew-Xold|/max|X| %.2e; rhs rel diff %.2e\n",
           trial, cond_boost, cn, ok, cudaGetErrorString(e), co, cudaGetErrorString(e2), resid(X), resid(Xo), dX_ / nX, dxy);
  }
  return 0;
}
