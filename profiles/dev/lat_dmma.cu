#include <cstdio>
__device__ __forceinline__ void dmma884(double (&d)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(d[0]), "+d"(d[1]) : "d"(a), "d"(b));
}
__global__ void k(double *out, long long *cyc, double x0) {
  double d[2] = {x0, x0 + 1}, a = 1.0000001 + threadIdx.x * 1e-9, b = 0.25;
  long long t0, t1;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) dmma884(d, a, b);
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
  double e[8][2];
#pragma unroll
  for (int q = 0; q < 8; q++) { e[q][0] = x0 + q; e[q][1] = x0 - q; }
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++)
#pragma unroll
    for (int q = 0; q < 8; q++) dmma884(e[q], a, b);
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
  double x = d[0] + d[1];
#pragma unroll
  for (int q = 0; q < 8; q++) x += e[q][0] + e[q][1];
  // shfl dependent
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31) + 1.0;
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
  // 16 independent dfma chains
  double c[16];
#pragma unroll
  for (int q = 0; q < 16; q++) c[q] = x + q;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++)
#pragma unroll
    for (int q = 0; q < 16; q++) c[q] = fma(c[q], a, b);
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
#pragma unroll
  for (int q = 0; q < 16; q++) x += c[q];
  __syncthreads();
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) __syncthreads();
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
  out[threadIdx.x] = x;
}
int main() {
  double *o; long long *c; cudaMalloc(&o, 8 * 1024); cudaMalloc(&c, 8 * 16);
  for (int nt : {32, 128, 256}) {
    k<<<1, nt>>>(o, c, 1.5); cudaDeviceSynchronize();
    long long h[16]; cudaMemcpy(h, c, 8 * 16, cudaMemcpyDeviceToHost);
    printf("threads %d: dep DMMA %.1f cyc; 8 indep DMMA chains: %.2f cyc/instr/warp; dep shfl64+dadd %.1f; 16-chain DFMA %.2f cyc/instr; syncthreads %.1f\n", nt, h[0] / 256.0, h[1] / 512.0, h[2] / 256.0, h[3] / 1024.0, h[4] / 64.0);
  }
}
