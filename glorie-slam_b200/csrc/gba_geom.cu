// gba_geom.cu -- geometry helpers next to the BA hot path, all context-free
// (no workspace) and asynchronous on the given stream:
//  rank 1: frame_distance, the O(n^2) keyframe-proximity metric that runs before every global BA
//          (reference src/lib/droid_kernels.cu:518-657,1441-1463; callers src/depth_video.py:166-196,
//          src/factor_graph.py:323-462);
//  rank 2: depth_filter, the multi-view consistency count behind the DSPO valid-depth mask
//          (src/lib/droid_kernels.cu:661-775,1494-1518; caller src/depth_video.py:326-361), and
//          align_scale_shift, the per-keyframe weighted scale/shift fit of the mono prior
//          (src/utils/common.py:402-438; callers src/depth_video.py:233,308);
//  rank 4: iproj and projmap, the rest of the reference module's reprojection surface
//          (src/lib/droid_kernels.cu:779-850,1520-1545 and :427-516,1466-1492; caller src/neural_point.py:156).
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
      float qq[4];
      rel_se3(poses + 7 * (size_t)i, poses + 7 * (size_t)j, qq, t);
      quat_to_rot(qq, R);
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

// ---------------------------------------------------------------------------------------------
// depth_filter
// ---------------------------------------------------------------------------------------------
#define DF_NT 256

// One thread per pixel of a listed keyframe; it visits the 6 neighbouring keyframes itself and keeps
// the count in a register.  The reference launches one CTA per (keyframe, neighbour, pixel block) and
// bumps the count with a global atomicAdd; here the counter is written once, needs no zero-fill and
// the source disparity is read once instead of six times.  Arithmetic follows the reference step by
// step (quaternion sandwich, true divisions, agreement test in double) because the result is a
// thresholded decision: a different rounding would flip borderline pixels.
__global__ void __launch_bounds__(DF_NT) k_depth_filter(const float *__restrict__ poses, const float *__restrict__ disps,
                                                        const float *__restrict__ intr,
                                                        const int64_t *__restrict__ inds,
                                                        const float *__restrict__ thresh, int n_frames, int ht, int wd,
                                                        float *__restrict__ counter) {
  __shared__ float s_rel[6][8];  // t[3], q[4], valid
  const int HW = ht * wd, b = blockIdx.y, tid = threadIdx.x;
  const int ix = (int)inds[b];
  if (tid < 6) {
    const int jx = (tid < 3) ? ix - tid - 1 : ix + tid;  // droid_kernels.cu:696
    const bool ok = jx >= 0 && jx < n_frames;
    s_rel[tid][7] = ok ? 1.f : 0.f;
    if (ok) {
      float qq[4], tt[3];
      rel_se3(poses + 7 * (size_t)ix, poses + 7 * (size_t)jx, qq, tt);
      s_rel[tid][0] = tt[0]; s_rel[tid][1] = tt[1]; s_rel[tid][2] = tt[2];
      s_rel[tid][3] = qq[0]; s_rel[tid][4] = qq[1]; s_rel[tid][5] = qq[2]; s_rel[tid][6] = qq[3];
    }
  }
  __syncthreads();
  const int k = blockIdx.x * DF_NT + tid;
  if (k >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int i = k / wd, j = k - i * wd;
  const float X0 = ((float)j - cx) / fx, X1 = ((float)i - cy) / fy, d = disps[(size_t)ix * HW + k];
  const double t = (double)thresh[b];
  int count = 0;
#pragma unroll 1
  for (int nb = 0; nb < 6; nb++) {
    if (s_rel[nb][7] == 0.f) continue;
    const int jx = (nb < 3) ? ix - nb - 1 : ix + nb;
    const float qx = s_rel[nb][3], qy = s_rel[nb][4], qz = s_rel[nb][5], qw = s_rel[nb][6];
    // Y = X + w * uv + q x uv,  uv = 2 q x X  with X = (X0, X1, 1)   (actSO3, lietorch so3 sandwich)
    const float uvx = 2.f * (qy * 1.f - qz * X1), uvy = 2.f * (qz * X0 - qx * 1.f), uvz = 2.f * (qx * X1 - qy * X0);
    float Yx = X0 + qw * uvx + (qy * uvz - qz * uvy);
    float Yy = X1 + qw * uvy + (qz * uvx - qx * uvz);
    float Yz = 1.f + qw * uvz + (qx * uvy - qy * uvx);
    Yx += d * s_rel[nb][0]; Yy += d * s_rel[nb][1]; Yz += d * s_rel[nb][2];
    const float uj = fx * (Yx / Yz) + cx, vj = fy * (Yy / Yz) + cy, dj = d / Yz;
    if (!(uj > -1e8f && uj < 1e8f && vj > -1e8f && vj < 1e8f)) continue;
    const int u0 = (int)floorf(uj), v0 = (int)floorf(vj);
    if (u0 >= 0 && v0 >= 0 && u0 < wd - 1 && v0 < ht - 1) {
      const float *dp = disps + (size_t)jx * HW + v0 * wd + u0;
      const float d00 = dp[0], d01 = dp[1], d10 = dp[wd], d11 = dp[wd + 1];
      const double inv = 1.0 / (double)dj;
      if (fabs(inv - 1.0 / (double)d00) < t || fabs(inv - 1.0 / (double)d01) < t ||
          fabs(inv - 1.0 / (double)d10) < t || fabs(inv - 1.0 / (double)d11) < t)
        count++;
    }
  }
  counter[(size_t)b * HW + k] = (float)count;
}

extern "C" int gba_depth_filter(const float *poses, const float *disps, const float *intrinsics, const int64_t *inds,
                                const float *thresh, int n_frames, int n_inds, int ht, int wd, float *counter,
                                void *stream) {
  if (n_inds == 0) return GBA_OK;  // empty index list: nothing to write (torch hands out NULL for empty tensors)
  if (!poses || !disps || !intrinsics || !inds || !thresh || !counter) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_depth_filter: NULL array argument");
    return GBA_EINVAL;
  }
  if (n_frames < 1 || n_inds < 0 || ht < 2 || wd < 2 || n_inds > 65535) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_depth_filter: bad sizes (n_inds <= 65535, ht, wd >= 2)");
    return GBA_EINVAL;
  }
  if (n_inds == 0) return GBA_OK;
  const dim3 grid((ht * wd + DF_NT - 1) / DF_NT, n_inds);
  k_depth_filter<<<grid, DF_NT, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, inds, thresh, n_frames, ht, wd,
                                                           counter);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_depth_filter: %s", cudaGetErrorString(e));
    return GBA_ECUDA;
  }
  return GBA_OK;
}

// ---------------------------------------------------------------------------------------------
// align_scale_shift
// ---------------------------------------------------------------------------------------------
#define AS_NT 256

template <int NS>
__device__ __forceinline__ void block_sum(double (&v)[NS], double (*s_red)[NS]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < NS; c++)
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v[c] += __shfl_xor_sync(0xffffffffu, v[c], o);
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int c = 0; c < NS; c++) s_red[warp][c] = v[c];
  __syncthreads();
#pragma unroll
  for (int c = 0; c < NS; c++) {
    double a = 0.0;
#pragma unroll
    for (int w = 0; w < AS_NT / 32; w++) a += s_red[w][c];
    v[c] = a;
  }
}

// One CTA per keyframe: five weighted sums, the 2x2 solve, then the weighted mean |residual| of the
// fit from the same (L1/L2-hot) rows.  The reference makes ~15 torch reductions over full-size
// temporaries for this; float32 partial sums per thread, combined in double.
__global__ void __launch_bounds__(AS_NT) k_align_scale_shift(const float *__restrict__ pred,
                                                             const float *__restrict__ target,
                                                             const float *__restrict__ weights, int HW,
                                                             float *__restrict__ scale, float *__restrict__ shift,
                                                             float *__restrict__ err) {
  __shared__ double s_red[AS_NT / 32][5];
  const size_t base = (size_t)blockIdx.x * HW;
  const float *p = pred + base, *t = target + base, *w = weights + base;
  float a[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  for (int k = threadIdx.x; k < HW; k += AS_NT) {
    const float pk = p[k], tk = t[k], wk = w[k], wp = wk * pk;
    a[0] = fmaf(wp, pk, a[0]); a[1] += wp; a[2] += wk; a[3] = fmaf(wp, tk, a[3]); a[4] = fmaf(wk, tk, a[4]);
  }
  double v[5];
#pragma unroll
  for (int c = 0; c < 5; c++) v[c] = (double)a[c];
  block_sum<5>(v, s_red);
  const double det = v[0] * v[2] - v[1] * v[1];
  const float s = (float)((v[2] * v[3] - v[1] * v[4]) / det), q = (float)((-v[1] * v[3] + v[0] * v[4]) / det);
  float e = 0.f;
  for (int k = threadIdx.x; k < HW; k += AS_NT) e = fmaf(fabsf(fmaf(s, p[k], q) - t[k]), w[k], e);
  double ev[5] = {(double)e, 0.0, 0.0, 0.0, 0.0};
  block_sum<5>(ev, s_red);
  if (threadIdx.x == 0) {
    scale[blockIdx.x] = s;
    shift[blockIdx.x] = q;
    err[blockIdx.x] = (float)(ev[0] / v[2]);
  }
}

extern "C" int gba_align_scale_shift(const float *prediction, const float *target, const float *weights, int n, int ht,
                                     int wd, float *scale, float *shift, float *avg_error, void *stream) {
  if (n == 0) return GBA_OK;
  if (!prediction || !target || !weights || !scale || !shift || !avg_error) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_align_scale_shift: NULL array argument");
    return GBA_EINVAL;
  }
  if (n < 0 || ht < 1 || wd < 1) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_align_scale_shift: bad sizes");
    return GBA_EINVAL;
  }
  if (n == 0) return GBA_OK;
  k_align_scale_shift<<<n, AS_NT, 0, (cudaStream_t)stream>>>(prediction, target, weights, ht * wd, scale, shift,
                                                             avg_error);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_align_scale_shift: %s", cudaGetErrorString(e));
    return GBA_ECUDA;
  }
  return GBA_OK;
}

// ---------------------------------------------------------------------------------------------
// iproj / projmap
// ---------------------------------------------------------------------------------------------
#define PJ_NT 256

// q * (x, y, 1) with the quaternion sandwich of the reference (actSO3, droid_kernels.cu:58-68)
__device__ __forceinline__ void rot_xy1(const float *q, float x, float y, float &X, float &Y, float &Z) {
  const float uvx = 2.f * (q[1] - q[2] * y), uvy = 2.f * (q[2] * x - q[0]), uvz = 2.f * (q[0] * y - q[1] * x);
  X = x + q[3] * uvx + (q[1] * uvz - q[2] * uvy);
  Y = y + q[3] * uvy + (q[2] * uvx - q[0] * uvz);
  Z = 1.f + q[3] * uvz + (q[0] * uvy - q[1] * uvx);
}

// Thread per pixel, grid (pixel blocks, keyframes): points = (R (x, y, 1) + t d) / d.  The three
// components of a pixel are stored by the thread that computed them (12 B stride; the reference does
// the same, there is nothing to reuse).
__global__ void __launch_bounds__(PJ_NT) k_iproj(const float *__restrict__ poses, const float *__restrict__ disps,
                                                 const float *__restrict__ intr, int HW, int wd,
                                                 float *__restrict__ points) {
  const int b = blockIdx.y, k = blockIdx.x * PJ_NT + threadIdx.x;
  if (k >= HW) return;
  const float *P = poses + 7 * (size_t)b;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int i = k / wd, j = k - i * wd;
  const float d = disps[(size_t)b * HW + k];
  float X, Y, Z;
  rot_xy1(P + 3, ((float)j - cx) / fx, ((float)i - cy) / fy, X, Y, Z);
  float *o = points + ((size_t)b * HW + k) * 3;
  o[0] = (X + d * P[0]) / d;
  o[1] = (Y + d * P[1]) / d;
  o[2] = (Z + d * P[2]) / d;
}

__global__ void __launch_bounds__(PJ_NT) k_projmap(const float *__restrict__ poses, const float *__restrict__ disps,
                                                   const float *__restrict__ intr, const int64_t *__restrict__ ii,
                                                   const int64_t *__restrict__ jj, int HW, int wd,
                                                   float *__restrict__ coords, float *__restrict__ valid) {
  __shared__ float s_rel[8];
  const int n = blockIdx.y, ix = (int)ii[n], jx = (int)jj[n];
  if (threadIdx.x == 0) {  // relSE3 (droid_kernels.cu:96-107)
    float qq[4], tt[3];
    rel_se3(poses + 7 * (size_t)ix, poses + 7 * (size_t)jx, qq, tt);
    s_rel[0] = tt[0]; s_rel[1] = tt[1]; s_rel[2] = tt[2];
    s_rel[3] = qq[0]; s_rel[4] = qq[1]; s_rel[5] = qq[2]; s_rel[6] = qq[3];
  }
  __syncthreads();
  const int k = blockIdx.x * PJ_NT + threadIdx.x;
  if (k >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int i = k / wd, j = k - i * wd;
  const float u = (float)j, v = (float)i, d = disps[(size_t)ix * HW + k];
  float X, Y, Z;
  rot_xy1(s_rel + 3, (u - cx) / fx, (v - cy) / fy, X, Y, Z);
  X += d * s_rel[0]; Y += d * s_rel[1]; Z += d * s_rel[2];
  float *c = coords + ((size_t)n * HW + k) * 3;
  const bool front = Z > 0.01f;
  c[0] = front ? fx * (X / Z) + cx : u;
  c[1] = front ? fy * (Y / Z) + cy : v;
  c[2] = 0.f;
  valid[(size_t)n * HW + k] = (Z > 0.25f) ? 1.0f : 0.0f;
}

static int geom_launch_check(const char *what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_create_err, sizeof(g_create_err), "%s: %s", what, cudaGetErrorString(e));
    return GBA_ECUDA;
  }
  return GBA_OK;
}

extern "C" int gba_iproj(const float *poses, const float *disps, const float *intrinsics, int n_frames, int ht, int wd,
                         float *points, void *stream) {
  if (n_frames == 0) return GBA_OK;
  if (!poses || !disps || !intrinsics || !points || n_frames < 0 || n_frames > 65535 || ht < 1 || wd < 1) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_iproj: NULL array argument or bad sizes (n_frames <= 65535)");
    return GBA_EINVAL;
  }
  k_iproj<<<dim3((ht * wd + PJ_NT - 1) / PJ_NT, n_frames), PJ_NT, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics,
                                                                                          ht * wd, wd, points);
  return geom_launch_check("gba_iproj");
}

extern "C" int gba_projmap(const float *poses, const float *disps, const float *intrinsics, const int64_t *ii,
                           const int64_t *jj, int n_pairs, int ht, int wd, float *coords, float *valid, void *stream) {
  if (n_pairs == 0) return GBA_OK;
  if (!poses || !disps || !intrinsics || !ii || !jj || !coords || !valid || n_pairs < 0 || n_pairs > 65535 || ht < 1 ||
      wd < 1) {
    snprintf(g_create_err, sizeof(g_create_err), "gba_projmap: NULL array argument or bad sizes (n_pairs <= 65535)");
    return GBA_EINVAL;
  }
  k_projmap<<<dim3((ht * wd + PJ_NT - 1) / PJ_NT, n_pairs), PJ_NT, 0, (cudaStream_t)stream>>>(
      poses, disps, intrinsics, ii, jj, ht * wd, wd, coords, valid);
  return geom_launch_check("gba_projmap");
}
