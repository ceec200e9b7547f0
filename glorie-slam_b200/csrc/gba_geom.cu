// gba_geom.cu -- geometry helpers next to the BA hot path:
// frame_distance, the O(n^2) keyframe-proximity metric that runs before every global BA
// (reference src/lib/droid_kernels.cu:518-657,1441-1463; callers src/depth_video.py:166-196,
// src/factor_graph.py:323-462).  Context-free: no workspace, asynchronous on the given stream.
#include <cuda_runtime.h>
#include <stdio.h>

#include "../../include/gba.h"
#include "gba_common.cuh"

extern thread_local char g_create_err[512];

#define FD_NT 256
#define FD_PB 8  // consecutive pairs per CTA: a meshgrid row keeps the source keyframe, its disparities stay in registers

// Mean induced flow of pair (i -> j): beta * |flow(G_ij)| + (1 - beta) * |flow(translation only)|,
// over pixels that land in front of the target camera (Z > 0.25); 1000 when < 75 % do.
// The reference spends one CTA per pair and three 256-wide shared-memory tree reductions; here a
// CTA walks FD_PB pairs, keeps the source disparities in registers while the source keyframe does
// not change, and reduces two sums per pair with warp shuffles.
template <int NV>
__global__ void __launch_bounds__(FD_NT) k_frame_distance(const float *__restrict__ poses,
                                                          const float *__restrict__ disps,
                                                          const float *__restrict__ intr,
                                                          const int64_t *__restrict__ ii,
                                                          const int64_t *__restrict__ jj, int n_pairs, int HW, int wd,
                                                          float beta, float *__restrict__ dist) {
  __shared__ float s_red[FD_NT / 32][2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float z[NV][4], x0[NV], y0[NV], u0[NV], v0[NV];
  int cur_i = -1;
#pragma unroll
  for (int g = 0; g < NV; g++) {
    const int px = (g * FD_NT + tid) * 4;
    u0[g] = (float)(px % wd);
    v0[g] = (float)(px / wd);
    x0[g] = (u0[g] - cx) / fx;
    y0[g] = (v0[g] - cy) / fy;
  }
  const float ifx = 1.0f / fx;
  for (int q = 0; q < FD_PB; q++) {
    const int n = blockIdx.x * FD_PB + q;
    if (n >= n_pairs) break;
    const int i = (int)ii[n], j = (int)jj[n];
    if (i != cur_i) {  // block-uniform
#pragma unroll
      for (int g = 0; g < NV; g++) {
        const int px = (g * FD_NT + tid) * 4;
        float4 zz = make_float4(0.f, 0.f, 0.f, 0.f);
        if (px < HW) zz = *reinterpret_cast<const float4 *>(disps + (size_t)i * HW + px);
        z[g][0] = zz.x; z[g][1] = zz.y; z[g][2] = zz.z; z[g][3] = zz.w;
      }
      cur_i = i;
    }
    // G_ij = T_j * T_i^-1 (relSE3, droid_kernels.cu:96-107), every thread for itself
    float R[9], t[3];
    {
      const float *Pi = poses + 7 * (size_t)i, *Pj = poses + 7 * (size_t)j, *qi = Pi + 3, *qj = Pj + 3;
      float qq[4];
      qq[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
      qq[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
      qq[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
      qq[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
      const float uv0 = 2.f * (qq[1] * Pi[2] - qq[2] * Pi[1]);
      const float uv1 = 2.f * (qq[2] * Pi[0] - qq[0] * Pi[2]);
      const float uv2 = 2.f * (qq[0] * Pi[1] - qq[1] * Pi[0]);
      t[0] = Pj[0] - (Pi[0] + qq[3] * uv0 + (qq[1] * uv2 - qq[2] * uv1));
      t[1] = Pj[1] - (Pi[1] + qq[3] * uv1 + (qq[2] * uv0 - qq[0] * uv2));
      t[2] = Pj[2] - (Pi[2] + qq[3] * uv2 + (qq[0] * uv1 - qq[1] * uv0));
      const float x = qq[0], y = qq[1], zq = qq[2], w = qq[3];
      R[0] = 1.f - 2.f * (y * y + zq * zq); R[1] = 2.f * (x * y - w * zq); R[2] = 2.f * (x * zq + w * y);
      R[3] = 2.f * (x * y + w * zq); R[4] = 1.f - 2.f * (x * x + zq * zq); R[5] = 2.f * (y * zq - w * x);
      R[6] = 2.f * (x * zq - w * y); R[7] = 2.f * (y * zq + w * x); R[8] = 1.f - 2.f * (x * x + y * y);
    }
    float accum = 0.f, valid = 0.f;
#pragma unroll
    for (int g = 0; g < NV; g++) {
      if ((g * FD_NT + tid) * 4 >= HW) continue;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const float xx = fmaf((float)k, ifx, x0[g]), yy = y0[g], zz = z[g][k];
        const float uu = u0[g] + (float)k, vv = v0[g];
        // full relative motion (droid_kernels.cu:586-604)
        {
          const float X = fmaf(R[0], xx, fmaf(R[1], yy, fmaf(t[0], zz, R[2])));
          const float Y = fmaf(R[3], xx, fmaf(R[4], yy, fmaf(t[1], zz, R[5])));
          const float Z = fmaf(R[6], xx, fmaf(R[7], yy, fmaf(t[2], zz, R[8])));
          const float iz = __fdividef(1.0f, Z);
          const float du = fmaf(fx * X, iz, cx) - uu, dv = fmaf(fy * Y, iz, cy) - vv;
          if (Z > 0.25f) {
            accum = fmaf(beta, sqrtf(fmaf(du, du, dv * dv)), accum);
            valid += beta;
          }
        }
        // translation only (droid_kernels.cu:606-628)
        {
          const float X = fmaf(zz, t[0], xx), Y = fmaf(zz, t[1], yy), Z = fmaf(zz, t[2], 1.0f);
          const float iz = __fdividef(1.0f, Z);
          const float du = fmaf(fx * X, iz, cx) - uu, dv = fmaf(fy * Y, iz, cy) - vv;
          if (Z > 0.25f) {
            accum = fmaf(1.0f - beta, sqrtf(fmaf(du, du, dv * dv)), accum);
            valid += 1.0f - beta;
          }
        }
      }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      accum += __shfl_xor_sync(0xffffffffu, accum, o);
      valid += __shfl_xor_sync(0xffffffffu, valid, o);
    }
    __syncthreads();
    if (lane == 0) {
      s_red[warp][0] = accum;
      s_red[warp][1] = valid;
    }
    __syncthreads();
    if (tid == 0) {
      float a = 0.f, v = 0.f;
#pragma unroll
      for (int w = 0; w < FD_NT / 32; w++) {
        a += s_red[w][0];
        v += s_red[w][1];
      }
      const float total = (float)HW;  // beta + (1 - beta) per pixel
      dist[n] = (v / (total + 1e-8f) < 0.75f) ? 1000.0f : a / v;
    }
  }
}

static int fail_geom(int code, const char *msg) {
  snprintf(g_create_err, sizeof(g_create_err), "gba_frame_distance: %s", msg);
  return code;
}

extern "C" int gba_frame_distance(const float *poses, const float *disps, const float *intrinsics, const int64_t *ii,
                                  const int64_t *jj, int n_pairs, int ht, int wd, float beta, float *dist,
                                  void *stream) {
  if (!poses || !disps || !intrinsics || !ii || !jj || !dist) return fail_geom(GBA_EINVAL, "NULL array argument");
  if (n_pairs < 0 || ht < 1 || wd < 1 || wd % 4 != 0) return fail_geom(GBA_EINVAL, "bad sizes (wd must be a multiple of 4)");
  if (n_pairs == 0) return GBA_OK;
  const int HW = ht * wd, grid = (n_pairs + FD_PB - 1) / FD_PB;
  cudaStream_t s = (cudaStream_t)stream;
  const int per_thread4 = (HW + FD_NT * 4 - 1) / (FD_NT * 4);  // float4 groups per thread
  if (per_thread4 <= 3) k_frame_distance<3><<<grid, FD_NT, 0, s>>>(poses, disps, intrinsics, ii, jj, n_pairs, HW, wd, beta, dist);
  else if (per_thread4 <= 5) k_frame_distance<5><<<grid, FD_NT, 0, s>>>(poses, disps, intrinsics, ii, jj, n_pairs, HW, wd, beta, dist);
  else if (per_thread4 <= 8) k_frame_distance<8><<<grid, FD_NT, 0, s>>>(poses, disps, intrinsics, ii, jj, n_pairs, HW, wd, beta, dist);
  else return fail_geom(GBA_ECAPACITY, "ht*wd larger than 8192 pixels is not supported");
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_geom(GBA_ECUDA, cudaGetErrorString(e));
  return GBA_OK;
}
