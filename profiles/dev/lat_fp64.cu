#include <cstdio>
__global__ void k(double *out, long long *cyc, double x0) {
  __shared__ double sm[64];
  double x = x0, y = 1.0000001;
  long long t0, t1;
  int lane = threadIdx.x & 31;
  // 1. dependent DFMA
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = fma(x, y, 1e-9);
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = (t1 - t0);
  // 2. dependent DMUL
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = x * y;
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = (t1 - t0);
  // 3. dependent shfl (64-bit)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = (t1 - t0);
  // 4. rcp.approx.ftz.f64 dependent
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); x = r; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = (t1 - t0);
  // 5. full reciprocal (seed + newton)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); double e = fma(-x, r, 1.0); x = fma(r, e, r); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = (t1 - t0);
  // 6. smem store -> syncthreads -> load (dependent)
  sm[threadIdx.x & 63] = x; __syncthreads();
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) { sm[(threadIdx.x + i) & 63] = x; __syncthreads(); x = sm[(threadIdx.x + i + 1) & 63] + 1.0; __syncthreads(); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[5] = (t1 - t0);
  // 7. independent DFMA throughput: 16 chains
  double c[16];
#pragma unroll
  for (int q = 0; q < 16; q++) c[q] = x + q;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++)
#pragma unroll
    for (int q = 0; q < 16; q++) c[q] = fma(c[q], y, 1e-9);
  t1 = clock64(); if (threadIdx.x == 0) cyc[6] = (t1 - t0);
#pragma unroll
  for (int q = 0; q < 16; q++) x += c[q];
  // 8. dependent FFMA
  float f = (float)x, g = 1.0000001f;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) f = fmaf(f, g, 1e-9f);
  t1 = clock64(); if (threadIdx.x == 0) cyc[7] = (t1 - t0);
  // 9. smem load dependent (pointer chase)
  __shared__ int idx[64];
  idx[threadIdx.x & 63] = (threadIdx.x * 7 + 1) & 63; __syncthreads();
  int p = lane;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) p = idx[p];
  t1 = clock64(); if (threadIdx.x == 0) cyc[8] = (t1 - t0);
  out[threadIdx.x] = x + f + p;
}
int main() {
  double *o; long long *c; cudaMalloc(&o, 8 * 256); cudaMalloc(&c, 8 * 16);
  for (int nt : {32, 256}) {
    k<<<1, nt>>>(o, c, 1.5); cudaDeviceSynchronize();
    long long h[16]; cudaMemcpy(h, c, 8 * 16, cudaMemcpyDeviceToHost);
    printf("threads %d: dep DFMA %.1f  dep DMUL %.1f  dep SHFL64 %.1f  rcp.approx.f64 %.1f  full rcp %.1f  sts+bar+lds+bar %.1f  16-chain DFMA per instr %.2f  dep FFMA %.1f  dep LDS %.1f\n", nt,
      h[0] / 256.0, h[1] / 256.0, h[2] / 256.0, h[3] / 256.0, h[4] / 256.0, h[5] / 64.0, h[6] / 1024.0, h[7] / 256.0, h[8] / 256.0);
  }
}
