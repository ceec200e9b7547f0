// gba_noalloc.cu -- gba_edge_blocks_static(): per-edge pose blocks without any allocation.
//
// Who this is for: code that is not allowed to call cudaMalloc at the point where it needs the
// per-edge Gauss-Newton blocks -- a real-time tracking thread that must not take the
// driver's allocation lock (cudaMalloc/cudaFree serialise with every other thread and can
// stall for milliseconds), a process whose whole GPU memory has already been handed to another
// allocator (PyTorch's caching allocator, a fixed arena) so that a late cudaMalloc would simply
// fail, or a plugin loaded into a host that forbids allocation after start-up.  The normal
// entry points (gba_create / gba_ba) take their workspaces from cudaMalloc; this one takes its
// scratch from a fixed array that is part of the library image.
//
// What it computes: for every factor-graph edge i -> j the 6x6 block H_jj = sum_pix w J_j^T J_j
// (upper triangle, 21 numbers) and the 6-vector v_j = sum_pix w J_j^T r -- the same numbers, in
// the same layout, that gba_ba produces internally and exposes through
// gba_get(GBA_GET_EDGE_HV).  The other three pose blocks of the edge follow from these with
// B = -Ad(G_ij)^T (H_ii = B H_jj B^T, H_ij = B H_jj, v_i = B v_j).
//
// Scratch: each edge is split over several CTAs (1024 pixels each); every CTA parks its 27
// partial sums in the static array, takes a ticket from a per-edge counter that also lives there,
// and the CTA that draws the last ticket adds the partials in a fixed order and writes the result.
// One kernel, no second pass, deterministic sums.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>

#include "../../include/gba.h"
#include "gba_common.cuh"

extern thread_local char g_create_err[512];  // the text gba_last_error(NULL) returns (gba_api.cu)

#define SB_NT 256
#define SB_PIX_PER_CTA (SB_NT * 4)
#define SB_MAX_CHUNKS ((GBA_STATIC_MAX_PIXELS + SB_PIX_PER_CTA - 1) / SB_PIX_PER_CTA)
#define SB_PARTIAL_FLOATS ((size_t)GBA_STATIC_MAX_EDGES * SB_MAX_CHUNKS * 32)

// The scratch itself: partial sums [edge][chunk][32] followed by one ticket counter per edge.
// About 2.6 MB, resident on every GPU that loads libgba.so until the process exits.
// NOTE for maintainers: in HOST code `g_static_scratch` is only a handle for the runtime, not a
// device address -- never pass it to a kernel or to cudaMemset directly.  Host code below asks
// the runtime for the real address once per device (cudaGetSymbolAddress) and uses that.
static __device__ __align__(256) float g_static_scratch[SB_PARTIAL_FLOATS + GBA_STATIC_MAX_EDGES];

__global__ void __launch_bounds__(SB_NT) k_edge_blocks_static(const float *__restrict__ poses,
                                                              const float *__restrict__ disps,
                                                              const float *__restrict__ intr,
                                                              const float *__restrict__ targets,
                                                              const float *__restrict__ weights,
                                                              const int64_t *__restrict__ ii,
                                                              const int64_t *__restrict__ jj, int n_frames, int HW, int wd,
                                                              int mode, float *scratch, float *__restrict__ out) {
  __shared__ float s_R[9], s_t[3];
  __shared__ float s_red[SB_NT / 32][28];
  __shared__ int s_last, s_stereo;
  const int e = blockIdx.y, chunk = blockIdx.x, nchunk = gridDim.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long i64 = ii[e], j64 = jj[e];
  // an edge that names a keyframe outside [0, n_frames) contributes nothing (its block stays zero)
  const bool in_range = i64 >= 0 && i64 < n_frames && j64 >= 0 && j64 < n_frames;
  const int i = in_range ? (int)i64 : 0, j = in_range ? (int)j64 : 0;
  float *partial = scratch;
  unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + SB_PARTIAL_FLOATS);

  if (tid == 0) {  // G_ij = T_j * T_i^-1 (droid_kernels.cu:96-107); stereo constant for i == j (:219-229)
    float q[4], t[3];
    if (i == j) {
      q[0] = q[1] = q[2] = 0.f; q[3] = 1.f;
      t[0] = -0.1f; t[1] = t[2] = 0.f;
    } else {
      rel_se3(poses + 7 * (size_t)i, poses + 7 * (size_t)j, q, t);
    }
    quat_to_rot(q, s_R);
    s_t[0] = t[0]; s_t[1] = t[1]; s_t[2] = t[2];
    s_stereo = (i == j) || !in_range;
  }
  __syncthreads();

  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float acc[32];
#pragma unroll
  for (int l = 0; l < 32; l++) acc[l] = 0.f;
  const int px = chunk * SB_PIX_PER_CTA + tid * 4;
  if (px < HW && !s_stereo) {  // self edges carry no pose terms (droid_kernels.cu:323,356)
    const float4 z4 = *reinterpret_cast<const float4 *>(disps + (size_t)i * HW + px);
    const float *tg = targets + (size_t)e * 2 * HW, *wg = weights + (size_t)e * 2 * HW;
    const float4 tu = ldg_stream4(tg + px), tv = ldg_stream4(tg + HW + px);
    const float4 wu = ldg_stream4(wg + px), wv = ldg_stream4(wg + HW + px);
    const float za[4] = {z4.x, z4.y, z4.z, z4.w};
    const float tua[4] = {tu.x, tu.y, tu.z, tu.w}, tva[4] = {tv.x, tv.y, tv.z, tv.w};
    const float wua[4] = {wu.x, wu.y, wu.z, wu.w}, wva[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int q = px + k;
      const float x0 = ((float)(q % wd) - cx) / fx, y0 = ((float)(q / wd) - cy) / fy;
      PixLin L;
      pix_linearize(s_R, s_t, fx, fy, cx, cy, mode, x0, y0, za[k], tua[k], tva[k], wua[k], wva[k], L);
      const float a0 = L.wu * L.ju0, a2 = L.wu * L.ju2, a3 = L.wu * L.ju3, a4 = L.wu * L.ju4, a5 = L.wu * L.ju5;
      const float b1 = L.wv * L.jv1, b2 = L.wv * L.jv2, b3 = L.wv * L.jv3, b4 = L.wv * L.jv4, b5 = L.wv * L.jv5;
      acc[0] += a0 * L.ju0; acc[2] += a0 * L.ju2; acc[3] += a0 * L.ju3; acc[4] += a0 * L.ju4; acc[5] += a0 * L.ju5;
      acc[6] += b1 * L.jv1; acc[7] += b1 * L.jv2; acc[8] += b1 * L.jv3; acc[9] += b1 * L.jv4; acc[10] += b1 * L.jv5;
      acc[11] += a2 * L.ju2 + b2 * L.jv2; acc[12] += a2 * L.ju3 + b2 * L.jv3; acc[13] += a2 * L.ju4 + b2 * L.jv4;
      acc[14] += a2 * L.ju5 + b2 * L.jv5; acc[15] += a3 * L.ju3 + b3 * L.jv3; acc[16] += a3 * L.ju4 + b3 * L.jv4;
      acc[17] += a3 * L.ju5 + b3 * L.jv5; acc[18] += a4 * L.ju4 + b4 * L.jv4; acc[19] += a4 * L.ju5 + b4 * L.jv5;
      acc[20] += a5 * L.ju5 + b5 * L.jv5;
      acc[21] += a0 * L.ru; acc[22] += b1 * L.rv; acc[23] += a2 * L.ru + b2 * L.rv; acc[24] += a3 * L.ru + b3 * L.rv;
      acc[25] += a4 * L.ru + b4 * L.rv; acc[26] += a5 * L.ru + b5 * L.rv;
    }
  }
  const float tot = warp_transpose_reduce32(acc, lane);
  if (lane < 28) s_red[warp][lane] = tot;
  __syncthreads();
  if (tid < 27) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < SB_NT / 32; w++) s += s_red[w][tid];
    partial[((size_t)e * nchunk + chunk) * 32 + tid] = s;
  }
  // publish the partials, then draw a ticket; the CTA holding the last one finishes the edge
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(&ticket[e], 1u) == (unsigned)(nchunk - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (tid < 32) {
    float s = 0.f;
    if (tid < 27)
      for (int c = 0; c < nchunk; c++) s += __ldcg(&partial[((size_t)e * nchunk + c) * 32 + tid]);
    out[(size_t)e * 32 + tid] = s;
  }
  if (tid == 0) ticket[e] = 0u;  // leave the counter clean for the next call as well
}

namespace {
std::mutex g_setup_mutex;
float *g_scratch_addr[64] = {nullptr};  // device address of g_static_scratch, per device ordinal

int fail(int code, const char *msg) {
  snprintf(g_create_err, sizeof(g_create_err), "gba_edge_blocks_static: %s", msg);
  return code;
}
}  // namespace

extern "C" int gba_edge_blocks_static(const float *poses, const float *disps, const float *intrinsics,
                                      const float *targets, const float *weights, const int64_t *ii,
                                      const int64_t *jj, int n_frames, int n_edges, int ht, int wd, int mode,
                                      float *out, void *stream) {
  if (!poses || !disps || !intrinsics || !targets || !weights || !ii || !jj || !out)
    return fail(GBA_EINVAL, "NULL array argument");
  if (n_frames < 1 || n_edges < 1 || ht < 1 || wd < 1) return fail(GBA_EINVAL, "sizes must be positive");
  if (mode != GBA_MODE_DROID_CUDA && mode != GBA_MODE_GEOM_PY) return fail(GBA_EINVAL, "unknown mode");
  const long long hw = (long long)ht * wd;
  if (hw % 4 != 0) return fail(GBA_EINVAL, "ht*wd must be a multiple of 4");
  // the kernel reads disps / targets / weights with 16-byte vector loads
  if ((((uintptr_t)disps | (uintptr_t)targets | (uintptr_t)weights) & 15) != 0)
    return fail(GBA_EINVAL, "disps, targets and weights must be 16-byte aligned");
  if (n_edges > GBA_STATIC_MAX_EDGES || hw > GBA_STATIC_MAX_PIXELS) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "%d edges x %lld pixels exceeds the built-in scratch (at most %d edges and %d pixels per keyframe); "
             "split the edge list or use gba_ba",
             n_edges, hw, GBA_STATIC_MAX_EDGES, GBA_STATIC_MAX_PIXELS);
    return fail(GBA_ECAPACITY, msg);
  }
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return fail(GBA_ECUDA, "no current CUDA device (this library has no CPU path)");
  }
  float *scratch;
  {
    // one-time set-up per device, with that device current: translate the symbol to its address
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[dev]) {
      float *p = nullptr;
      err = cudaGetSymbolAddress((void **)&p, g_static_scratch);
      if (err != cudaSuccess || !p) {
        cudaGetLastError();
        return fail(GBA_ECUDA, cudaGetErrorString(err));
      }
      g_scratch_addr[dev] = p;
    }
    scratch = g_scratch_addr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  // the array still holds whatever the previous call left; the ticket counters must start at zero
  err = cudaMemsetAsync(scratch + SB_PARTIAL_FLOATS, 0, (size_t)n_edges * sizeof(unsigned int), s);
  if (err != cudaSuccess) return fail(GBA_ECUDA, cudaGetErrorString(err));
  const int nchunk = (int)((hw + SB_PIX_PER_CTA - 1) / SB_PIX_PER_CTA);
  k_edge_blocks_static<<<dim3(nchunk, n_edges), SB_NT, 0, s>>>(poses, disps, intrinsics, targets, weights, ii, jj,
                                                               n_frames, (int)hw, wd, mode, scratch, out);
  err = cudaGetLastError();
  if (err != cudaSuccess) return fail(GBA_ECUDA, cudaGetErrorString(err));
  return GBA_OK;
}
