// gba_select.cu -- the host-side loops around the BA hot path that GlORIE-SLAM runs in Python, as kernels:
//   * proximity edge selection with non-maximum suppression (reference src/factor_graph.py:386-462,
//     FactorGraph.add_backend_proximity_factors: meshgrid -> video.distance -> thresholds -> local-window edges ->
//     sort -> greedy suppression with one .item() per candidate),
//   * the valid-depth mask after depth_filter (src/depth_video.py:349-361: count >= visible_num, per-keyframe
//     nanmedian of the surviving depths, depth < 3 median),
//   * the mono_thres edge / eta filter before DSPO (src/depth_video.py:245-261).
// Integer / boolean outputs: bit-exact against the reference's own functions (tests/golden/prox.npz).
// SURVEY.md section 8(f) ranks 1 and 2.  No gba_ctx; scratch is passed in by the caller.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/gba.h"

extern thread_local char g_create_err[512];

namespace {
int sel_fail(int code, const char *msg) {
  snprintf(g_create_err, 512, "%s", msg);
  return code;
}

#define SEL_NT 1024

// exclusive block scan of one int per thread (1024 threads); returns the exclusive prefix, *total = block sum
__device__ __forceinline__ int block_scan_excl(int v, int *s_warp, int *total) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += u;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  if (w == 0) {
    int y = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, y, o);
      if (lane >= o) y += u;
    }
    s_warp[lane] = y;
  }
  __syncthreads();
  const int excl = x - v + (w ? s_warp[w - 1] : 0);
  *total = s_warp[31];
  __syncthreads();
  return excl;
}

struct ProxParams {
  const float *d;          // [ilen * jlen] bidirectional frame distance, row = i - t_loop, column = j - t_start
  float *dm;               // [ilen * jlen] scratch: the masked / suppressed copy the selection works on
  unsigned *key[2];        // [n] scratch (radix sort ping-pong): float bits of the candidate distances
  int *val[2];             // [n] scratch: flat pair index of the candidates
  int *out_ii, *out_jj;    // [cap] edge list
  int *out_n;              // [0] edges written, [1] status: 0 ok, 1 = the reference returns 0 (too few edges), 2 = capacity
  int t_start, t_loop, t_end, ilen, jlen, nms, radius, max_factors, loop, cap;
  float thresh;
};

// ONE CTA.  Every step keeps the reference's order of operations (factor_graph.py:399-456).
__global__ void __launch_bounds__(SEL_NT) k_prox_select(ProxParams p) {
  __shared__ int s_warp[32];
  __shared__ int s_cnt[32][256 + 1];  // per warp, per digit
  __shared__ int s_base[256], s_run[256], s_tile[256];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.ilen * p.jlen;
  const float INF = __int_as_float(0x7f800000);
  // (1) d[ii - radius < jj] = inf; d[d > thresh] = inf            (:402-403)
  for (int k = tid; k < n; k += SEL_NT) {
    const int i = p.t_loop + k / p.jlen, j = p.t_start + k % p.jlen;
    float v = p.d[k];
    if (i - p.radius < j) v = INF;
    if (v > p.thresh) v = INF;
    p.dm[k] = v;
  }
  __syncthreads();
  // (2) edges of the local window [i - radius - 1, i), both directions; their cells leave the candidates (:406-413).
  // A column index j - t_start < 0 wraps around like the reference's negative tensor index does.
  int n_es = 0;
  {
    // first edge slot of keyframe i: 2 * sum_{i' in [t_loop, i)} min(radius + 1, i')
    for (int i = p.t_loop + tid; i < p.t_end; i += SEL_NT) {
      long long off = 0;
      for (int q = p.t_loop; q < i; q++) off += min(p.radius + 1, q);
      int e = (int)(2 * off);
      for (int j = max(i - p.radius - 1, 0); j < i; j++) {
        if (e + 1 < p.cap) {
          p.out_ii[e] = i; p.out_jj[e] = j;
          p.out_ii[e + 1] = j; p.out_jj[e + 1] = i;
        }
        e += 2;
        int dj = j - p.t_start;
        if (dj < 0) dj += p.jlen;
        if (dj >= 0 && dj < p.jlen) p.dm[(i - p.t_loop) * p.jlen + dj] = INF;
      }
    }
    long long tot = 0;
    for (int q = p.t_loop; q < p.t_end; q++) tot += min(p.radius + 1, q);
    n_es = (int)(2 * tot);
  }
  __syncthreads();
  // (3) candidates = cells that are still <= thresh, in ascending flat order (ordered compaction)
  int ncand = 0;
  for (int base = 0; base < n; base += SEL_NT) {
    const int k = base + tid;
    const float v = k < n ? p.dm[k] : INF;
    const int keep = v <= p.thresh ? 1 : 0;
    int tot;
    const int pos = block_scan_excl(keep, s_warp, &tot);
    if (keep) {
      p.key[0][ncand + pos] = __float_as_uint(v);  // distances are >= 0: their bit patterns order like the floats
      p.val[0][ncand + pos] = k;
    }
    ncand += tot;
  }
  __syncthreads();
  // (4) stable LSD radix sort by distance, 4 x 8 bits: ties keep ascending flat index (torch.sort, :416)
  int cur = 0;
  for (int pass = 0; pass < 4; pass++) {
    const int shift = 8 * pass;
    for (int i = tid; i < 256; i += SEL_NT) s_base[i] = 0;
    __syncthreads();
    for (int k = tid; k < ncand; k += SEL_NT) atomicAdd(&s_base[(p.key[cur][k] >> shift) & 255], 1);
    __syncthreads();
    if (tid < 256) {  // exclusive scan of the 256 digit counts (8 warps)
      const int v = s_base[tid];
      int x = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += u;
      }
      if (lane == 31) s_warp[warp] = x;
      s_run[tid] = x - v;  // exclusive inside the warp
    }
    __syncthreads();
    if (tid < 256) {
      int add = 0;
      for (int w = 0; w < warp; w++) add += s_warp[w];
      s_base[tid] = s_run[tid] + add;
    }
    __syncthreads();
    if (tid < 256) s_run[tid] = 0;
    __syncthreads();
    for (int base = 0; base < ncand; base += SEL_NT) {
      const int k = base + tid;
      const bool ok = k < ncand;
      const unsigned kv = ok ? p.key[cur][k] : 0u;
      const int vv = ok ? p.val[cur][k] : 0;
      const int dg = ok ? (int)((kv >> shift) & 255u) : 256;  // 256: padding lanes, a digit of their own
      for (int i = lane; i < 257; i += 32) s_cnt[warp][i] = 0;
      __syncwarp();
      const unsigned peers = __match_any_sync(0xffffffffu, dg);
      const int rank = __popc(peers & ((1u << lane) - 1u));
      if (rank == 0) s_cnt[warp][dg] = __popc(peers);
      __syncthreads();
      // per digit: exclusive offsets of the warps inside this tile, and the tile total
      if (tid < 256) {
        int acc = 0;
        for (int w = 0; w < 32; w++) {
          const int c = s_cnt[w][tid];
          s_cnt[w][tid] = acc;
          acc += c;
        }
        s_tile[tid] = acc;
      }
      __syncthreads();
      if (ok) {
        const int dst = s_base[dg] + s_run[dg] + s_cnt[warp][dg] + rank;
        p.key[cur ^ 1][dst] = kv;
        p.val[cur ^ 1][dst] = vv;
      }
      __syncthreads();
      if (tid < 256) s_run[tid] += s_tile[tid];
      __syncthreads();
    }
    cur ^= 1;
    __syncthreads();
  }
  // (5) greedy selection in sorted order with suppression of the (2 nms + 1)^2 neighbourhood (:421-452), warp 0,
  // 32 candidates at a time: a candidate is alive while its cell has not been suppressed by an earlier selection
  if (warp == 0) {
    const int *sv = p.val[cur];
    int loop_edges = 0;
    bool stop = false, overflow = false;
    for (int base = 0; base < ncand && !stop; base += 32) {
      const int c = base + lane < ncand ? sv[base + lane] : -1;
      bool alive = c >= 0 && !(p.dm[c] > p.thresh);
      unsigned mask = __ballot_sync(0xffffffffu, alive);
      while (mask && !stop) {
        const int L = __ffs(mask) - 1;
        const int k = __shfl_sync(0xffffffffu, c, L);
        if (n_es > p.max_factors) {  // :426-427, tested before every selection
          stop = true;
          break;
        }
        const int di = k / p.jlen, dj = k % p.jlen, i = p.t_loop + di, j = p.t_start + dj;
        if (p.loop) {  // :434-445: the still-close pairs around (i, j) that span more than 20 keyframes
          const int si = i - 1 + lane / 3, sj = j - 1 + lane % 3;
          bool add = lane < 9 && si >= max(i - 1, p.t_loop) && si < min(i + 2, p.t_end) && sj >= max(j - 1, p.t_start) &&
                     sj < min(j + 2, p.t_end);
          if (add) add = p.d[(si - p.t_loop) * p.jlen + (sj - p.t_start)] <= p.thresh && si != sj && si - sj > 20;
          const unsigned am = __ballot_sync(0xffffffffu, add);
          if (add) {
            const int e = n_es + __popc(am & ((1u << lane) - 1u));
            if (e < p.cap) {
              p.out_ii[e] = si;
              p.out_jj[e] = sj;
            } else
              overflow = true;
          }
          n_es += __popc(am);
          loop_edges += __popc(am);
        } else {  // :446-448
          if (lane == 0) {
            if (n_es + 1 < p.cap) {
              p.out_ii[n_es] = i; p.out_jj[n_es] = j;
              p.out_ii[n_es + 1] = j; p.out_jj[n_es + 1] = i;
            } else
              overflow = true;
          }
          n_es += 2;
        }
        // :450 d[di-nms : di+nms+1, dj-nms : dj+nms+1] = inf (clipped)
        const int side = 2 * p.nms + 1;
        for (int q = lane; q < side * side; q += 32) {
          const int ri = di - p.nms + q / side, rj = dj - p.nms + q % side;
          if (ri >= 0 && ri < p.ilen && rj >= 0 && rj < p.jlen) p.dm[ri * p.jlen + rj] = INF;
        }
        __syncwarp();
        alive = alive && lane > L && !(p.dm[c] > p.thresh);
        mask = __ballot_sync(0xffffffffu, alive);
      }
    }
    overflow = __any_sync(0xffffffffu, overflow);
    if (lane == 0) {
      p.out_n[0] = min(n_es, p.cap);
      p.out_n[1] = overflow ? 2 : ((n_es < 3 || (p.loop && loop_edges == 0)) ? 1 : 0);  // :455-456
    }
  }
}
}  // namespace

namespace {
// ---------------------------------------------------------------------------------------------------
// Valid-depth mask (src/depth_video.py:340-361): depth = 1 / disp; pixels with count < visible_num are NaN;
// median = torch.nanmedian over the keyframe (the LOWER middle value of the non-NaN depths); mask = depth < 3 median.
// One CTA per keyframe: exact k-th smallest by a 4 x 8-bit radix select over the keyframe's pixels (L2 resident).
// By-products for the mono_thres filter: the number of valid pixels and the mean disparity of the keyframe.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned order_key(float v) {  // monotone map float -> uint (negatives included)
  const unsigned b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void __launch_bounds__(256) k_valid_depth_mask(const float *disps, const float *count, int hw, float visible_num,
                                                          unsigned char *mask, int *valid_cnt, float *mean_disp) {
  __shared__ int s_hist[256];
  __shared__ int s_sel[2];  // chosen digit, remaining rank
  __shared__ double s_sum[8];
  __shared__ int s_m;
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float *z = disps + (size_t)f * hw, *cn = count + (size_t)f * hw;
  // number of multi-view consistent pixels, mean disparity (fp64 accumulation)
  int m = 0;
  double acc = 0.0;
  for (int px = tid; px < hw; px += 256) {
    m += (cn[px] >= visible_num && !isnan(1.0f / z[px])) ? 1 : 0;
    acc += (double)z[px];
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    m += __shfl_xor_sync(0xffffffffu, m, o);
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
  }
  if (tid == 0) s_m = 0;
  __syncthreads();
  if (lane == 0) {
    atomicAdd(&s_m, m);
    s_sum[warp] = acc;
  }
  __syncthreads();
  m = s_m;
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; w++) t += s_sum[w];
    mean_disp[f] = (float)(t / (double)hw);
  }
  float med = __int_as_float(0x7fc00000);  // NaN when no pixel survives (nanmedian of all-NaN)
  if (m > 0) {
    int rank = (m - 1) / 2;  // lower median, 0-based
    unsigned prefix = 0, pmask = 0;
    for (int pass = 3; pass >= 0; pass--) {
      const int shift = 8 * pass;
      s_hist[tid] = 0;
      __syncthreads();
      for (int px = tid; px < hw; px += 256) {
        const float dpt = 1.0f / z[px];
        if (cn[px] >= visible_num && !isnan(dpt)) {
          const unsigned k = order_key(dpt);
          if ((k & pmask) == prefix) atomicAdd(&s_hist[(k >> shift) & 255], 1);
        }
      }
      __syncthreads();
      if (tid == 0) {
        int r = rank, dgt = 0;
        for (; dgt < 256; dgt++) {
          if (r < s_hist[dgt]) break;
          r -= s_hist[dgt];
        }
        s_sel[0] = dgt;
        s_sel[1] = r;
      }
      __syncthreads();
      prefix |= (unsigned)s_sel[0] << shift;
      pmask |= 255u << shift;
      rank = s_sel[1];
      __syncthreads();
    }
    const unsigned b = (prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix;
    med = __uint_as_float(b);
  }
  const float lim = 3.0f * med;
  for (int px = tid; px < hw; px += 256) {
    const float dpt = 1.0f / z[px];
    mask[(size_t)f * hw + px] = (cn[px] >= visible_num && dpt < lim) ? 1 : 0;
  }
  // valid_d.sum() of the reference counts the MASK, not the multi-view test
  __syncthreads();
  int vc = 0;
  for (int px = tid; px < hw; px += 256) vc += (cn[px] >= visible_num && (1.0f / z[px]) < lim) ? 1 : 0;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) vc += __shfl_xor_sync(0xffffffffu, vc, o);
  if (tid == 0) s_m = 0;
  __syncthreads();
  if (lane == 0) atomicAdd(&s_m, vc);
  __syncthreads();
  if (tid == 0) valid_cnt[f] = s_m;
}

// mono_thres edge filter (src/depth_video.py:245-261).  invalid[f] for the first n keyframes, keep[e] for every edge,
// src_kept[f] = keyframe f is still the source of a kept edge (the caller masks eta with it: eta rows follow
// unique(ii)).  One CTA.
__global__ void __launch_bounds__(1024) k_mono_filter(const float *error, const float *avg_disp, const float *scale,
                                                      const int *valid_cnt, int n, int hw, float thres, const int64_t *ii,
                                                      const int64_t *jj, int E, int n_frames, unsigned char *invalid,
                                                      unsigned char *keep, unsigned char *src_kept) {
  const int tid = threadIdx.x;
  for (int f = tid; f < n_frames; f += 1024) {
    src_kept[f] = 0;
    if (f < n) {
      const float e = error[f];
      invalid[f] = ((e / avg_disp[f] > thres) || isnan(e) || (scale[f] < 0.f) || ((float)valid_cnt[f] < (float)hw * 0.5f)) ? 1 : 0;
    }
  }
  __syncthreads();
  for (int e = tid; e < E; e += 1024) {
    const long long i = ii[e], j = jj[e];
    const bool bad = (i >= 0 && i < n && invalid[i]) || (j >= 0 && j < n && invalid[j]);
    keep[e] = bad ? 0 : 1;
    if (!bad && i >= 0 && i < n_frames) src_kept[i] = 1;
  }
}
}  // namespace

extern "C" size_t gba_proximity_scratch_bytes(int ilen, int jlen) {
  const size_t n = (size_t)(ilen > 0 ? ilen : 0) * (size_t)(jlen > 0 ? jlen : 0);
  return 20 * n + 256;
}

extern "C" int gba_proximity_select(const float *d, int t_start, int t_start_loop, int t_end, int nms, int radius,
                                    float thresh, int max_factors, int loop, void *scratch, size_t scratch_bytes,
                                    int *out_ii, int *out_jj, int cap, int *out_n, void *stream) {
  if (!d || !scratch || !out_ii || !out_jj || !out_n) return sel_fail(GBA_EINVAL, "gba_proximity_select: NULL argument");
  if (!loop) t_start_loop = t_start;  // factor_graph.py:387-388
  if (t_start < 0 || t_start_loop < t_start || t_end <= t_start_loop || nms < 0 || radius < 0 || cap < 2)
    return sel_fail(GBA_EINVAL, "gba_proximity_select: bad window / parameters");
  ProxParams p;
  p.ilen = t_end - t_start_loop;
  p.jlen = t_end - t_start;
  const size_t n = (size_t)p.ilen * p.jlen;
  if (n > (size_t)1 << 30) return sel_fail(GBA_ECAPACITY, "gba_proximity_select: too many keyframe pairs");
  if (scratch_bytes < gba_proximity_scratch_bytes(p.ilen, p.jlen)) return sel_fail(GBA_EINVAL, "gba_proximity_select: scratch too small");
  char *s = (char *)scratch;
  p.d = d;
  p.dm = (float *)s;
  p.key[0] = (unsigned *)(s + 4 * n);
  p.key[1] = (unsigned *)(s + 8 * n);
  p.val[0] = (int *)(s + 12 * n);
  p.val[1] = (int *)(s + 16 * n);
  p.out_ii = out_ii; p.out_jj = out_jj; p.out_n = out_n;
  p.t_start = t_start; p.t_loop = t_start_loop; p.t_end = t_end; p.nms = nms; p.radius = radius;
  p.max_factors = max_factors; p.loop = loop ? 1 : 0; p.cap = cap; p.thresh = thresh;
  k_prox_select<<<1, SEL_NT, 0, (cudaStream_t)stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return sel_fail(GBA_ECUDA, cudaGetErrorString(e));
  return GBA_OK;
}

extern "C" int gba_valid_depth_mask(const float *disps, const float *count, int n, int ht, int wd, int visible_num,
                                    uint8_t *mask, int *valid_cnt, float *mean_disp, void *stream) {
  if (!disps || !count || !mask || !valid_cnt || !mean_disp) return sel_fail(GBA_EINVAL, "gba_valid_depth_mask: NULL argument");
  if (n < 0 || ht < 1 || wd < 1) return sel_fail(GBA_EINVAL, "gba_valid_depth_mask: bad sizes");
  if (n == 0) return GBA_OK;
  k_valid_depth_mask<<<n, 256, 0, (cudaStream_t)stream>>>(disps, count, ht * wd, (float)visible_num, mask, valid_cnt, mean_disp);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return sel_fail(GBA_ECUDA, cudaGetErrorString(e));
  return GBA_OK;
}

extern "C" int gba_mono_filter(const float *error, const float *avg_disp, const float *scale, const int *valid_cnt, int n,
                               int ht, int wd, float mono_thres, const int64_t *ii, const int64_t *jj, int n_edges,
                               int n_frames, uint8_t *invalid, uint8_t *keep, uint8_t *src_kept, void *stream) {
  if (!error || !avg_disp || !scale || !valid_cnt || !invalid || !src_kept || (n_edges > 0 && (!ii || !jj || !keep)))
    return sel_fail(GBA_EINVAL, "gba_mono_filter: NULL argument");
  if (n < 0 || n > n_frames || n_edges < 0) return sel_fail(GBA_EINVAL, "gba_mono_filter: bad sizes");
  k_mono_filter<<<1, 1024, 0, (cudaStream_t)stream>>>(error, avg_disp, scale, valid_cnt, n, ht * wd, mono_thres, ii, jj,
                                                       n_edges, n_frames, invalid, keep, src_kept);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return sel_fail(GBA_ECUDA, cudaGetErrorString(e));
  return GBA_OK;
}
