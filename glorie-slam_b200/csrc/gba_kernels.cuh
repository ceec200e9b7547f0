// gba_kernels.cuh -- index build and the per-edge / per-frame kernels of the dense BA.
#pragma once
#include "gba_common.cuh"

// ============================================================================================
// Index build (once per call, on device).  Replaces the host-side work of the reference:
// torch::_unique over cat(ts, ii) (droid_kernels.cu:1337-1345) and the D2H + argsort + CSR loop
// that every accum_cuda call repeats (droid_kernels.cu:948-984).
// ============================================================================================

// one CTA: depth-keyframe list kx (sorted unique of [t0,t1) U ii), frame -> k map
__global__ void __launch_bounds__(1024) k_prep_frames(BaParams p, const int64_t *kx_global, int n_kx, int first_pass) {
  __shared__ int s_carry;
  __shared__ int s_warp[32];
  const int tid = threadIdx.x;
  if (tid == 0 && first_pass) {  // a sharded begin builds the CSR twice: errors of the first pass must survive
    p.st->err = 0;
    p.st->chol_fail = 0;
    p.st->last_fail = 0;
  }
  for (int f = tid; f < p.N; f += blockDim.x) p.fmap[f] = 0;
  __syncthreads();
  if (kx_global) {
    for (int k = tid; k < n_kx; k += blockDim.x) {
      const long long f = kx_global[k];
      if (f < 0 || f >= p.N) atomicOr(&p.st->err, 1);
      else {
        p.kx[k] = (int)f;
        p.fmap[f] = k + 1;
      }
    }
    __syncthreads();
    for (int n = tid; n < p.E; n += blockDim.x) {
      const long long i = p.ii[n];
      if (i < 0 || i >= p.N) atomicOr(&p.st->err, 1);
      else if (p.fmap[i] == 0) atomicOr(&p.st->err, 4);
    }
    __syncthreads();
    for (int f = tid; f < p.N; f += blockDim.x) p.fmap[f] -= 1;
    for (int k = tid; k < n_kx; k += blockDim.x) p.deg[k] = 0;
    if (tid == 0) p.st->K = n_kx;
    return;
  }
  if (p.mode == 0)
    for (int f = p.t0 + tid; f < p.t1 && f < p.N; f += blockDim.x)
      if (f >= 0) p.fmap[f] = 1;
  for (int n = tid; n < p.E; n += blockDim.x) {
    const long long i = p.ii[n];
    if (i < 0 || i >= p.N) atomicOr(&p.st->err, 1);
    else p.fmap[i] = 1;
  }
  if (tid == 0) s_carry = 0;
  __syncthreads();
  // exclusive scan of the presence flags, 1024 frames per round
  for (int base = 0; base < p.N; base += blockDim.x) {
    const int f = base + tid;
    const int flag = (f < p.N) ? p.fmap[f] : 0;
    int v = flag;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if ((tid & 31) >= o) v += u;
    }
    if ((tid & 31) == 31) s_warp[tid >> 5] = v;
    __syncthreads();
    if (tid < 32) {
      int w = s_warp[tid];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, w, o);
        if (tid >= o) w += u;
      }
      s_warp[tid] = w;
    }
    __syncthreads();
    const int incl = v + ((tid >> 5) ? s_warp[(tid >> 5) - 1] : 0) + s_carry;
    if (f < p.N) {
      if (flag) {
        p.kx[incl - 1] = f;
        p.fmap[f] = incl - 1;
        p.deg[incl - 1] = 0;
      } else
        p.fmap[f] = -1;
    }
    __syncthreads();
    if (tid == blockDim.x - 1) s_carry = incl;
    __syncthreads();
  }
  if (tid == 0) p.st->K = s_carry;
}

// edge -> k, out-degree histogram, range checks
__global__ void k_prep_edges(BaParams p) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= p.E) return;
  const long long i = p.ii[n], j = p.jj[n];
  if (i < 0 || i >= p.N || j < 0 || j >= p.N) {
    atomicOr(&p.st->err, 1);
    p.ek[n] = -1;
    return;
  }
  const int k = p.fmap[i];
  p.ek[n] = k;
  if (k >= 0) atomicAdd(&p.deg[k], 1);
}

// one CTA: CSR row pointers, per-frame Gram offsets, max degree
__global__ void __launch_bounds__(1024) k_prep_scan(BaParams p) {
  __shared__ int s_warp[32];
  __shared__ int s_carry, s_maxdeg, s_ncc;
  __shared__ unsigned long long s_gcarry;
  __shared__ unsigned long long s_gwarp[32];
  const int tid = threadIdx.x, K = p.st->K;
  if (tid == 0) {
    s_carry = 0;
    s_maxdeg = 0;
    s_ncc = 0;
    s_gcarry = 0ull;
    p.eptr[0] = 0;
  }
  __syncthreads();
  for (int base = 0; base < K; base += blockDim.x) {
    const int k = base + tid;
    const int d = (k < K) ? p.deg[k] : 0;
    const int nb = (6 * d + 1 + 7) / 8;
    const unsigned long long g = (k < K && d > 0) ? (unsigned long long)(nb * 8) * (nb * 8) : 0ull;
    int v = d;
    unsigned long long gv = g;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      const unsigned long long gu = __shfl_up_sync(0xffffffffu, gv, o);
      if ((tid & 31) >= o) {
        v += u;
        gv += gu;
      }
    }
    if ((tid & 31) == 31) {
      s_warp[tid >> 5] = v;
      s_gwarp[tid >> 5] = gv;
    }
    atomicMax(&s_maxdeg, d);
    if (k < K && d > p.tc_maxdeg && (!p.owned || p.owned[k])) p.cc_list[atomicAdd(&s_ncc, 1)] = k;
    __syncthreads();
    if (tid < 32) {
      int w = s_warp[tid];
      unsigned long long gw = s_gwarp[tid];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, w, o);
        const unsigned long long gu = __shfl_up_sync(0xffffffffu, gw, o);
        if (tid >= o) {
          w += u;
          gw += gu;
        }
      }
      s_warp[tid] = w;
      s_gwarp[tid] = gw;
    }
    __syncthreads();
    const int incl = v + ((tid >> 5) ? s_warp[(tid >> 5) - 1] : 0) + s_carry;
    const unsigned long long gincl = gv + ((tid >> 5) ? s_gwarp[(tid >> 5) - 1] : 0ull) + s_gcarry;
    if (k < K) {
      p.eptr[k + 1] = incl;
      p.goff[k] = gincl - g;
    }
    __syncthreads();
    if (tid == blockDim.x - 1) {
      s_carry = incl;
      s_gcarry = gincl;
    }
    __syncthreads();
  }
  if (tid == 0) {
    p.st->max_deg = s_maxdeg;
    p.st->n_cc = s_ncc;
    p.st->gram_floats = s_gcarry;
  }
}

// one warp per depth keyframe: stable fill of its edge list (edge order, like argsort's CSR)
__global__ void k_prep_fill(BaParams p) {
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (k >= p.st->K) return;
  int cnt = p.eptr[k];
  // the scan is latency bound (one L2 load per 32 edges): keep 8 loads in flight
  for (int base = 0; base < p.E; base += 256) {
    int ekv[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int n = base + 32 * u + lane;
      ekv[u] = (n < p.E) ? p.ek[n] : -1;
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const bool m = ekv[u] == k;
      const unsigned b = __ballot_sync(0xffffffffu, m);
      if (m) p.eidx[cnt + __popc(b & ((1u << lane) - 1u))] = base + 32 * u + lane;
      cnt += __popc(b);
    }
  }
}

// ============================================================================================
// Per-iteration kernels
// ============================================================================================

// Relative pose of every edge from the current poses (droid_kernels.cu:96-107 relSE3).
__global__ void k_edge_geo(BaParams p) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= p.E) return;
  const int i = (int)p.ii[n], j = (int)p.jj[n];
  EdgeGeo g;
  float q[4], t[3];
  g.flags = 0;
  if (i == j) {  // stereo pair: droid_kernels.cu:219-229
    q[0] = q[1] = q[2] = 0.f;
    q[3] = 1.f;
    t[0] = -0.1f;
    t[1] = t[2] = 0.f;
    g.flags = 1;
  } else {
    rel_se3(p.poses + 7 * (size_t)i, p.poses + 7 * (size_t)j, q, t);
  }
  quat_to_rot(q, g.R);
  g.t[0] = t[0];
  g.t[1] = t[1];
  g.t[2] = t[2];
  const int pi = i - p.t0, pj = j - p.t0;
  g.pi = (pi >= 0 && pi < p.P) ? pi : -1;  // update_lhs drops negative indices (droid_kernels.cu:1146)
  g.pj = (pj >= 0 && pj < p.P) ? pj : -1;
  g.pad = 0;
  p.geo[n] = g;
}

// --------------------------------------------------------------------------------------------
// Pass A: linearisation.  CTA = (pixel chunk, depth keyframe k); each thread owns NV groups of 4
// consecutive pixels and walks the out-edges of k.  Reads target/weight once (float4, streaming),
// keeps the per-pixel depth diagonal C and w in registers, reduces the 27 entries of (H_jj, v_j)
// per edge with a 31-shuffle transposed warp reduction and a per-warp shared-memory stage
// (no global atomics).  Replaces projective_transform_kernel (droid_kernels.cu:176-424) and the
// accum_cuda calls for C and w (droid_kernels.cu:1399-1401).
// --------------------------------------------------------------------------------------------
#define LIN_EB 8  // edges per shared-memory batch

// occupancy targets measured at config E: 128 x 8 pixels at 4 CTAs per SM (125 registers) 0.32 ms, 3 CTAs (168
// registers) 0.38, 5 CTAs (96, spills) 0.35; 128 x 4 pixels: 6 CTAs (80 registers) 0.34, 4 CTAs 0.41
template <int NT, int NV, int MODE, bool IL>
__global__ void __launch_bounds__(NT, (NT >= 256) ? 2 : (NV == 1 ? 6 : 4)) k_linearize(BaParams p) {
  constexpr int NW = NT / 32;
  __shared__ EdgeGeo s_geo[LIN_EB];
  __shared__ int s_eid[LIN_EB];
  __shared__ float s_red[NW][LIN_EB][28];

  const int k = blockIdx.y;
  if (k >= p.st->K) return;
  if (p.owned && !p.owned[k]) return;
  const int chunk = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int frame = p.kx[k];
  const int HW = p.HW;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const PixConst pk = make_pix_const(p.intr);

  // wd is a multiple of 4 (checked in gba_create): the 4 pixels of a group share one image row
  int px[NV];
  float x0[NV], y0[NV], z[NV][4], cacc[NV][4], wacc[NV][4];
#pragma unroll
  for (int g = 0; g < NV; g++) {
    px[g] = (chunk * NV + g) * (NT * 4) + tid * 4;
    float4 zz = make_float4(0.f, 0.f, 0.f, 0.f);
    if (px[g] < HW) zz = *reinterpret_cast<const float4 *>(p.disps + (size_t)frame * HW + px[g]);
    z[g][0] = zz.x; z[g][1] = zz.y; z[g][2] = zz.z; z[g][3] = zz.w;
    x0[g] = ((float)(px[g] % p.wd) - cx) / fx;  // droid_kernels.cu:290-291
    y0[g] = ((float)(px[g] / p.wd) - cy) / fy;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      cacc[g][i] = 0.f;
      wacc[g][i] = 0.f;
    }
  }

  const int e0 = p.eptr[k], e1 = p.eptr[k + 1];
  for (int eb = e0; eb < e1; eb += LIN_EB) {
    const int nb = min(LIN_EB, e1 - eb);
    __syncthreads();
    if (tid < nb) s_eid[tid] = p.eidx[eb + tid];
    __syncthreads();
    for (int i = tid; i < nb * 16; i += NT)
      reinterpret_cast<float *>(s_geo)[i] = reinterpret_cast<const float *>(p.geo + s_eid[i >> 4])[i & 15];
    __syncthreads();

    for (int q = 0; q < nb; q++) {
      const int n = s_eid[q];
      const float *R = s_geo[q].R, *t = s_geo[q].t;
      const bool stereo = s_geo[q].flags & 1;
      const float *tg = p.targets + (size_t)n * 2 * HW, *wg = p.weights + (size_t)n * 2 * HW;
      float4 tu[NV], tv[NV], wu[NV], wv[NV];
#pragma unroll
      for (int g = 0; g < NV; g++) {
        if (px[g] < HW) {
          ldg_uv4(tg, HW, px[g], IL, tu[g], tv[g]);
          ldg_uv4(wg, HW, px[g], IL, wu[g], wv[g]);
        } else {
          tu[g] = tv[g] = wu[g] = wv[g] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      // 27 sums of (H_jj, v_j), kept as 10 packed pairs + 7 scalars (FFMA2: two sums per issue slot)
      unsigned long long P[10];
      float s0 = 0.f, s6 = 0.f, s17 = 0.f, s20 = 0.f, s21 = 0.f, s22 = 0.f;
#pragma unroll
      for (int i = 0; i < 10; i++) P[i] = 0ull;
#pragma unroll
      for (int g = 0; g < NV; g++) {
        const float tua[4] = {tu[g].x, tu[g].y, tu[g].z, tu[g].w}, tva[4] = {tv[g].x, tv[g].y, tv[g].z, tv[g].w};
        const float wua[4] = {wu[g].x, wu[g].y, wu[g].z, wu[g].w}, wva[4] = {wv[g].x, wv[g].y, wv[g].z, wv[g].w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
          PixHat L;
          pix_hat<MODE>(R, t, pk, fmaf((float)i, pk.ifx, x0[g]), y0[g], z[g][i], tua[i], tva[i], wua[i], wva[i], L);
          // depth diagonal keeps the un-zeroed weight even for stereo pairs (droid_kernels.cu:320-323)
          const float qu = L.wu * L.jzu, qv = L.wv * L.jzv;
          cacc[g][i] = fmaf(qv, L.jzv, fmaf(qu, L.jzu, cacc[g][i]));
          wacc[g][i] = fmaf(qv, L.rv, fmaf(qu, L.ru, wacc[g][i]));
          if (stereo) continue;  // block-uniform: self edges carry no pose terms (droid_kernels.cu:323,356)
          // ju' = (e, 0, nae, nab, opa, nb), jv' = (0, e, nbe, nopb, ab, a)
          const unsigned long long U1 = pack2(L.nae, L.nab), U2 = pack2(L.opa, L.nb), U3 = pack2(L.nab, L.opa);
          const unsigned long long V1 = pack2(L.nbe, L.nopb), V2 = pack2(L.ab, L.a), V3 = pack2(L.nopb, L.ab);
          const unsigned long long WU = pack2(L.wu, L.wu), WV = pack2(L.wv, L.wv);
          const unsigned long long A23 = mul2(WU, U1), A45 = mul2(WU, U2), B23 = mul2(WV, V1), B45 = mul2(WV, V2);
          float a2, a3, a4, a5, b2, b3, b4, b5;
          unpack2(A23, a2, a3);
          unpack2(A45, a4, a5);
          unpack2(B23, b2, b3);
          unpack2(B45, b4, b5);
          const float a0 = L.wu * L.e, b1 = L.wv * L.e;
          const unsigned long long A0 = pack2(a0, a0), B1 = pack2(b1, b1);
          // H_jj upper triangle, row major: 0:(0,0) 1:(0,1)=0 2..5:(0,2..5) 6:(1,1) 7..10:(1,2..5)
          // 11..14:(2,2..5) 15..17:(3,3..5) 18,19:(4,4..5) 20:(5,5); v_j: 21..26
          s0 = fmaf(a0, L.e, s0);
          fma2(P[0], A0, U1);  // (0,2),(0,3)
          fma2(P[1], A0, U2);  // (0,4),(0,5)
          s6 = fmaf(b1, L.e, s6);
          fma2(P[2], B1, V1);  // (1,2),(1,3)
          fma2(P[3], B1, V2);  // (1,4),(1,5)
          fma2(P[4], pack2(a2, a2), U1);  // (2,2),(2,3)
          fma2(P[4], pack2(b2, b2), V1);
          fma2(P[5], pack2(a2, a2), U2);  // (2,4),(2,5)
          fma2(P[5], pack2(b2, b2), V2);
          fma2(P[6], pack2(a3, a3), U3);  // (3,3),(3,4)
          fma2(P[6], pack2(b3, b3), V3);
          s17 = fmaf(b3, L.a, fmaf(a3, L.nb, s17));  // (3,5)
          fma2(P[7], pack2(a4, a4), U2);  // (4,4),(4,5)
          fma2(P[7], pack2(b4, b4), V2);
          s20 = fmaf(b5, L.a, fmaf(a5, L.nb, s20));  // (5,5)
          // v_j (w r J = w' r' j')
          s21 = fmaf(a0, L.ru, s21);
          s22 = fmaf(b1, L.rv, s22);
          const unsigned long long RU = pack2(L.ru, L.ru), RV = pack2(L.rv, L.rv);
          fma2(P[8], RU, A23);
          fma2(P[8], RV, B23);
          fma2(P[9], RU, A45);
          fma2(P[9], RV, B45);
        }
      }
      float acc[32];
#pragma unroll
      for (int i = 27; i < 32; i++) acc[i] = 0.f;
      acc[0] = s0; acc[1] = 0.f; acc[6] = s6; acc[17] = s17; acc[20] = s20; acc[21] = s21; acc[22] = s22;
      unpack2(P[0], acc[2], acc[3]);
      unpack2(P[1], acc[4], acc[5]);
      unpack2(P[2], acc[7], acc[8]);
      unpack2(P[3], acc[9], acc[10]);
      unpack2(P[4], acc[11], acc[12]);
      unpack2(P[5], acc[13], acc[14]);
      unpack2(P[6], acc[15], acc[16]);
      unpack2(P[7], acc[18], acc[19]);
      unpack2(P[8], acc[23], acc[24]);
      unpack2(P[9], acc[25], acc[26]);
      const float tot = warp_transpose_reduce32(acc, lane);
      if (lane < 28) s_red[warp][q][lane] = tot;
    }
    __syncthreads();
    for (int i = tid; i < nb * 27; i += NT) {
      const int q = i / 27, l = i - q * 27;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < NW; w++) s += s_red[w][q][l];
      p.hpart[((size_t)s_eid[q] * p.nchunk + chunk) * 32 + l] = s;
    }
  }

  if (p.motion_only) return;
  // depth diagonal: C = sum Cii + m*alpha + (1-m)*eta, w = sum wi - m*alpha*(disp - sens)
  // (droid_kernels.cu:1397-1401; ba.py:92-103), Q = 1/C
  const float alpha = 0.05f;
#pragma unroll
  for (int g = 0; g < NV; g++) {
    if (px[g] >= HW) continue;
    const float4 sn = *reinterpret_cast<const float4 *>(p.sens + (size_t)frame * HW + px[g]);
    const float4 et = *reinterpret_cast<const float4 *>(p.eta + (size_t)k * HW + px[g]);
    const float sna[4] = {sn.x, sn.y, sn.z, sn.w}, eta4[4] = {et.x, et.y, et.z, et.w};
    float qo[4], wo[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float m = sna[i] > 0.f ? 1.f : 0.f;
      const float C = cacc[g][i] + m * alpha + (1.f - m) * eta4[i];
      wo[i] = wacc[g][i] - m * alpha * (z[g][i] - sna[i]);
      qo[i] = 1.0f / C;
    }
    *reinterpret_cast<float4 *>(p.Q + (size_t)k * HW + px[g]) = make_float4(qo[0], qo[1], qo[2], qo[3]);
    *reinterpret_cast<float4 *>(p.W + (size_t)k * HW + px[g]) = make_float4(wo[0], wo[1], wo[2], wo[3]);
  }
}

// --------------------------------------------------------------------------------------------
// Per edge: sum the chunk partials, derive the four pose blocks from H_jj through
// B = -Ad(G_ij)^T (H_ii = B H B^T, H_ij = B H, H_ji = H B^T, v_i = B v_j) and add them to the
// reduced camera system (fp64, lower triangle).  Replaces SparseBlock::update_lhs/update_rhs
// (droid_kernels.cu:1131-1173,1377-1384) -- no D2H copy, no triplets.
// --------------------------------------------------------------------------------------------
__global__ void k_edge_finish(BaParams p) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= p.E) return;
  if (p.ek[n] < 0 || (p.owned && !p.owned[p.ek[n]])) return;
  double hv[27];
  for (int l = 0; l < 27; l++) {
    float s = 0.f;
    for (int c = 0; c < p.nchunk; c++) s += p.hpart[((size_t)n * p.nchunk + c) * 32 + l];
    hv[l] = (double)s;
    p.edge_hv[(size_t)n * 32 + l] = s;
  }
  const EdgeGeo g = p.geo[n];
  if (g.pi < 0 && g.pj < 0) return;
  // positions in the (permuted) reduced system
  const int pi = g.pi >= 0 ? p.perm[g.pi] : -1, pj = g.pj >= 0 ? p.perm[g.pj] : -1;
  double H[6][6];
  {
    int l = 0;
    for (int r = 0; r < 6; r++)
      for (int c = r; c < 6; c++) {
        H[r][c] = hv[l];
        H[c][r] = hv[l];
        l++;
      }
  }
  const double *v = hv + 21;
  double *rhs = p.sys + (size_t)p.st->n_slots * GBA_NB2;
  // jj,jj block and v_j
  if (pj >= 0) {
    for (int r = 0; r < 6; r++) {
      for (int c = 0; c <= r; c++) atomicAdd(sys_at(p, 6 * pj + r, 6 * pj + c), H[r][c]);
      atomicAdd(rhs + 6 * pj + r, v[r]);
    }
  }
  if (pi < 0) return;
  // BH = B H (columns of H mapped through B)
  double BH[6][6];
  for (int c = 0; c < 6; c++) {
    double x[6], y[6];
    for (int r = 0; r < 6; r++) x[r] = H[r][c];
    apply_B<double>(g.R, g.t, x, y);
    for (int r = 0; r < 6; r++) BH[r][c] = y[r];
  }
  // ii,ii block = (B H) B^T: row r of BH mapped through B
  for (int r = 0; r < 6; r++) {
    double x[6], y[6];
    for (int c = 0; c < 6; c++) x[c] = BH[r][c];
    apply_B<double>(g.R, g.t, x, y);
    for (int c = 0; c <= r; c++) atomicAdd(sys_at(p, 6 * pi + r, 6 * pi + c), y[c]);
  }
  {
    double y[6];
    apply_B<double>(g.R, g.t, v, y);
    for (int r = 0; r < 6; r++) atomicAdd(rhs + 6 * pi + r, y[r]);
  }
  if (pj >= 0 && pi != pj) {
    // off-diagonal: block (pi,pj) = B H, block (pj,pi) = (B H)^T; store the lower one
    for (int r = 0; r < 6; r++)
      for (int c = 0; c < 6; c++) {
        if (pi > pj) atomicAdd(sys_at(p, 6 * pi + r, 6 * pj + c), BH[r][c]);
        else atomicAdd(sys_at(p, 6 * pj + c, 6 * pi + r), BH[r][c]);
      }
  }
}

// --------------------------------------------------------------------------------------------
// Pass B: Schur Gram.  For depth keyframe k with out-edges n_1..n_d the rows
// a_n[c][pix] = sqrt(Q) * sum_{u,v} w J_z J_j[c]  (c = 0..5) and the extra row sqrt(Q) * w_k are
// rebuilt tile by tile in shared memory (never written to HBM) and the Gram matrix
// G = A A^T is accumulated in registers, 8x8 outputs per thread over the block-upper triangle.
// G[6n+r, 6m+c] = sum_pix E_n[r] Q E_m[c];  G[6n+r, 6d] = sum_pix E_n[r] Q w.
// Replaces Eii/Eij materialisation + accum + EEt6x6/Ev6x1 (droid_kernels.cu:340-341,373-374,
// 1001-1093,1222-1296): the (i,j,k) triples of schur_block are exactly the block pairs of G.
// --------------------------------------------------------------------------------------------
#define GRAM_T 32   // pixels per shared-memory tile
#define GRAM_NT 256
#define GRAM_MAXCW 6  // at most 6 consumer warps (192 register blocks per pass), at least 2 producers
#define GRAM_PB 5     // edges whose inputs a producer lane keeps in flight

// pitch (floats) of one pixel row of the shared tile: >= 8 * nbs, a multiple of 4 (float4 reads)
// and == 4 (mod 32) so that the 32 producer lanes of a column store spread over 8 banks
__host__ __device__ __forceinline__ int gram_pitch(int nbs) { return ((8 * nbs + 27) / 32) * 32 + 4; }

__device__ __forceinline__ void named_bar_sync(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(GRAM_NT) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(GRAM_NT) : "memory"); }
// Warp-specialised: producer warps rebuild the sqrt(Q)-scaled E rows of a 32-pixel tile into one
// of two shared-memory buffers while consumer warps accumulate the previous tile into their 8x8
// register blocks (block-upper triangle of G = A A^T) with packed FFMA2.  FULL/EMPTY named
// barriers (ids 1-4) hand the buffers over; nobody waits at a CTA-wide barrier inside the pixel
// loop.  Diagonal blocks are enumerated first, so only the first consumer warp(s) also carry the
// extra column G[:, 6 deg] = sum_pix E Q w.
__global__ void __launch_bounds__(GRAM_NT, 2) k_gram(BaParams p) {
  extern __shared__ float s_tile[];
  const int k = p.cc_list[blockIdx.y];  // the keyframes with more than tc_maxdeg out-edges; the others are k_gram_tc's
  const int e0 = p.eptr[k], deg = p.eptr[k + 1] - e0;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int Rw = 6 * deg;              // rows of E; shared-memory column Rw holds sqrt(Q) * w
  const int nb = (Rw + 7) / 8;         // 8-row blocks of G proper
  const int nbs = (Rw + 1 + 7) / 8;    // 8-column chunks of a shared-memory row (incl. the w column)
  const int Rp = nbs * 8;              // pitch of G in the gram buffer (k_prep_scan uses the same)
  const int RS = gram_pitch(nbs), half = 4 * nbs;  // row = [cols 0-3 of all blocks | cols 4-7 of all blocks]
  const int nblk = nb * (nb + 1) / 2;
  const int ncw = min(GRAM_MAXCW, (nblk + 31) / 32);  // consumer warps
  const int npw = GRAM_NT / 32 - ncw;                 // producer warps
  const int npass = (nblk + ncw * 32 - 1) / (ncw * 32);
  const int HW = p.HW, frame = p.kx[k];
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const PixConst pk = make_pix_const(p.intr);
  const int ntile = (HW + GRAM_T - 1) / GRAM_T;
  const int nloc = (ntile - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles of this CTA
  float *G = p.gram + p.goff[k];
  const bool producer = warp >= ncw;

  // padding columns of both buffers stay zero; data columns are rewritten for every tile
  for (int i = tid; i < 2 * GRAM_T * RS; i += GRAM_NT) s_tile[i] = 0.f;
  // per-edge geometry and edge ids of this keyframe, staged once behind the two tile buffers
  EdgeGeo *s_geo = reinterpret_cast<EdgeGeo *>(s_tile + 2 * GRAM_T * RS);
  int *s_eid = reinterpret_cast<int *>(s_geo + deg);
  for (int i = tid; i < deg; i += GRAM_NT) s_eid[i] = p.eidx[e0 + i];
  for (int i = tid; i < deg * 16; i += GRAM_NT)
    reinterpret_cast<float *>(s_geo)[i] = reinterpret_cast<const float *>(p.geo + p.eidx[e0 + (i >> 4)])[i & 15];
  __syncthreads();

  for (int pass = 0; pass < npass; pass++) {
    if (producer) {
      const int pw = warp - ncw;
      const int nq = (deg - pw + npw - 1) / npw;  // edges of this warp: q = pw + j * npw
      const size_t fo = (size_t)frame * HW, ko = (size_t)k * HW;
      // one (edge, pixel) item: E row sqrt(Q) w' jz' j' into the half-split shared row
      auto emit = [&](float *row, int q, bool pv, float x0, float y0, float zq, float sq, float tu_, float tv_,
                      float wu_, float wv_) {
        const EdgeGeo &g = s_geo[q];
        float e[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (pv && !(g.flags & 1)) {
          PixHat L;
          if (p.mode == 0) pix_hat<0>(g.R, g.t, pk, x0, y0, zq, tu_, tv_, wu_, wv_, L);
          else pix_hat<1>(g.R, g.t, pk, x0, y0, zq, tu_, tv_, wu_, wv_, L);
          const float su = sq * L.wu * L.jzu, sv = sq * L.wv * L.jzv;
          e[0] = su * L.e;
          e[1] = sv * L.e;
          e[2] = fmaf(su, L.nae, sv * L.nbe);
          e[3] = fmaf(su, L.nab, sv * L.nopb);
          e[4] = fmaf(su, L.opa, sv * L.ab);
          e[5] = fmaf(su, L.nb, sv * L.a);
        }
        // half-split row: columns [0..3] of every 8-block first, then columns [4..7]; the three
        // column pairs (6q, 6q+2, 6q+4 are even) never straddle a half
#pragma unroll
        for (int c = 0; c < 6; c += 2) {
          const int col = 6 * q + c;
          *reinterpret_cast<float2 *>(row + ((col >> 2) & 1) * half + ((col >> 3) << 2) + (col & 3)) =
              make_float2(e[c], e[c + 1]);
        }
      };
      const int woff = ((Rw >> 2) & 1) * half + ((Rw >> 3) << 2) + (Rw & 3);
      if (nq <= GRAM_PB) {
        // common case: all edges of this warp fit one batch.  The producer is latency bound, so the
        // inputs of tile it+1 are requested before tile it is computed (register double buffer).
        const float *wgp[GRAM_PB];
#pragma unroll
        for (int u = 0; u < GRAM_PB; u++) {
          const int n = (u < nq) ? s_eid[pw + u * npw] : 0;
          wgp[u] = p.weights + (size_t)n * 2 * HW;
        }
        float tu_[GRAM_PB], tv_[GRAM_PB], wu_[GRAM_PB], wv_[GRAM_PB], zq = 1.f, qv = 0.f, wk = 0.f;
        auto fetch = [&](int it) {
          const int pix = (blockIdx.x + it * gridDim.x) * GRAM_T + lane;
          const bool pv = pix < HW;
          zq = pv ? p.disps[fo + pix] : 1.f;
          qv = pv ? p.Q[ko + pix] : 0.f;
          wk = (pw == 0 && pv) ? p.W[ko + pix] : 0.f;
#pragma unroll
          for (int u = 0; u < GRAM_PB; u++) {
            const bool ok = pv && u < nq;
            const int ou = p.il ? 2 * pix : pix, ov = p.il ? 2 * pix + 1 : HW + pix;
            tu_[u] = tv_[u] = 0.f;  // the E rows carry no residual: the targets are not read in this pass
            wu_[u] = ok ? ldg_stream(wgp[u] + ou) : 0.f;
            wv_[u] = ok ? ldg_stream(wgp[u] + ov) : 0.f;
          }
        };
        if (nloc > 0) fetch(0);
        for (int it = 0; it < nloc; it++) {
          const int b = it & 1;
          float *row = s_tile + b * GRAM_T * RS + lane * RS;
          const int pix = (blockIdx.x + it * gridDim.x) * GRAM_T + lane;
          const bool pv = pix < HW;
          const float x0 = ((float)(pix % p.wd) - cx) / fx, y0 = ((float)(pix / p.wd) - cy) / fy;
          // move the current inputs out of the way, then request the next tile's
          float ctu[GRAM_PB], ctv[GRAM_PB], cwu[GRAM_PB], cwv[GRAM_PB];
#pragma unroll
          for (int u = 0; u < GRAM_PB; u++) {
            ctu[u] = tu_[u]; ctv[u] = tv_[u]; cwu[u] = wu_[u]; cwv[u] = wv_[u];
          }
          const float czq = zq, csq = sqrtf(qv), cwk = wk;
          if (it + 1 < nloc) fetch(it + 1);
          if (it >= 2) named_bar_sync(3 + b);  // consumers are done with this buffer (tile it-2)
#pragma unroll
          for (int u = 0; u < GRAM_PB; u++)
            if (u < nq) emit(row, pw + u * npw, pv, x0, y0, czq, csq, ctu[u], ctv[u], cwu[u], cwv[u]);
          if (pw == 0) row[woff] = csq * cwk;
          __threadfence_block();
          named_bar_arrive(1 + b);  // buffer b is full
        }
      } else {
        for (int it = 0; it < nloc; it++) {
          const int b = it & 1;
          float *row = s_tile + b * GRAM_T * RS + lane * RS;
          if (it >= 2) named_bar_sync(3 + b);
          const int pix = (blockIdx.x + it * gridDim.x) * GRAM_T + lane;
          const bool pv = pix < HW;
          const float zq = pv ? p.disps[fo + pix] : 1.f;
          const float sq = pv ? sqrtf(p.Q[ko + pix]) : 0.f;
          const float x0 = ((float)(pix % p.wd) - cx) / fx, y0 = ((float)(pix / p.wd) - cy) / fy;
          const float wk = (pw == 0 && pv) ? p.W[ko + pix] : 0.f;
          for (int j0 = 0; j0 < nq; j0 += GRAM_PB) {
            float tu_[GRAM_PB], tv_[GRAM_PB], wu_[GRAM_PB], wv_[GRAM_PB];
#pragma unroll
            for (int u = 0; u < GRAM_PB; u++) {
              tu_[u] = tv_[u] = wu_[u] = wv_[u] = 0.f;
              if (j0 + u < nq && pv) {
                const int n = s_eid[pw + (j0 + u) * npw];
                const float *wg = p.weights + (size_t)n * 2 * HW;
                const int ou = p.il ? 2 * pix : pix, ov = p.il ? 2 * pix + 1 : HW + pix;
                wu_[u] = ldg_stream(wg + ou);
                wv_[u] = ldg_stream(wg + ov);
              }
            }
#pragma unroll
            for (int u = 0; u < GRAM_PB; u++)
              if (j0 + u < nq) emit(row, pw + (j0 + u) * npw, pv, x0, y0, zq, sq, tu_[u], tv_[u], wu_[u], wv_[u]);
          }
          if (pw == 0) row[woff] = sq * wk;
          __threadfence_block();
          named_bar_arrive(1 + b);
        }
      }
    } else {
      // block enumeration: the nb diagonal blocks first, then bi < bj row major
      const int blk = pass * ncw * 32 + tid;
      int bi = 0, bj = 0;
      const bool active = blk < nblk;
      const bool diag = blk < nb;
      if (diag) bi = bj = blk;
      else if (active) {
        int rem = blk - nb;
        while (rem >= nb - 1 - bi) {
          rem -= nb - 1 - bi;
          bi++;
        }
        bj = bi + 1 + rem;
      }
      const bool wdiag = (pass * ncw * 32 + warp * 32) < nb;  // warp-uniform: this warp owns diagonal blocks
      unsigned long long acc[8][4];
      float accw[8];
#pragma unroll
      for (int r = 0; r < 8; r++) {
        accw[r] = 0.f;
#pragma unroll
        for (int c = 0; c < 4; c++) acc[r][c] = 0ull;
      }
      for (int it = 0; it < nloc; it++) {
        const int b = it & 1;
        const float *pa = s_tile + b * GRAM_T * RS + 4 * bi;
        const float *pb = s_tile + b * GRAM_T * RS + 4 * bj;
        const float *pw_ = s_tile + b * GRAM_T * RS + ((Rw >> 2) & 1) * half + ((Rw >> 3) << 2) + (Rw & 3);
        named_bar_sync(1 + b);  // wait until the producers filled buffer b
        if (active) {
#pragma unroll 8
          for (int pp = 0; pp < GRAM_T; pp++) {
            const float4 a0 = *reinterpret_cast<const float4 *>(pa), a1 = *reinterpret_cast<const float4 *>(pa + half);
            const float4 b0 = *reinterpret_cast<const float4 *>(pb), b1 = *reinterpret_cast<const float4 *>(pb + half);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const unsigned long long bp[4] = {pack2(b0.x, b0.y), pack2(b0.z, b0.w), pack2(b1.x, b1.y), pack2(b1.z, b1.w)};
#pragma unroll
            for (int r = 0; r < 8; r++) {
              const unsigned long long ap = pack2(a[r], a[r]);
#pragma unroll
              for (int c = 0; c < 4; c++) fma2(acc[r][c], ap, bp[c]);
            }
            if (wdiag) {
              const float wq = *pw_;
#pragma unroll
              for (int r = 0; r < 8; r++) accw[r] = fmaf(a[r], wq, accw[r]);
            }
            pa += RS;
            pb += RS;
            pw_ += RS;
          }
        }
        if (it + 2 < nloc) named_bar_arrive(3 + b);  // buffer b may be refilled
      }
      if (active) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
          for (int c = 0; c < 4; c++) {
            float v0, v1;
            unpack2(acc[r][c], v0, v1);
            float *dst = G + (size_t)(8 * bi + r) * Rp + 8 * bj + 2 * c;
            if (8 * bj + 2 * c != Rw) {  // column Rw belongs to the E Q w sums below
              if (gridDim.x == 1) dst[0] = v0;
              else atomicAdd(dst, v0);
            }
            if (8 * bj + 2 * c + 1 != Rw) {
              if (gridDim.x == 1) dst[1] = v1;
              else atomicAdd(dst + 1, v1);
            }
          }
          if (diag && 8 * bi + r < Rw) {
            float *dst = G + (size_t)(8 * bi + r) * Rp + Rw;
            if (gridDim.x == 1) *dst = accw[r];
            else atomicAdd(dst, accw[r]);
          }
        }
      }
    }
    __syncthreads();  // both roles: all tiles of this pass consumed before the buffers are reused
  }
}

// --------------------------------------------------------------------------------------------
// Per depth keyframe: subtract its Schur contribution from the reduced camera system.
// Edge rows carry pose jj[n]; the keyframe's own pose row is E_i = sum_n B_n E_n, so its blocks
// follow algebraically: S[k,m] = sum_n B_n G[n,m], S[k,k] = sum_{n,m} B_n G[n,m] B_m^T.
// Replaces schur_block's host-side adjacency + O(P^2 deg^2) triple loop + sparse assembly
// (droid_kernels.cu:1241-1311) and `A - S` (droid_kernels.cu:1175-1177,1407).
// --------------------------------------------------------------------------------------------
#define SCHUR_NT 128

__device__ __forceinline__ float gram_at(const float *G, int Rp, int r, int c) {
  return (r <= c) ? G[(size_t)r * Rp + c] : G[(size_t)c * Rp + r];
}

__global__ void __launch_bounds__(SCHUR_NT) k_schur(BaParams p) {
  extern __shared__ float s_U[];  // [6][Rw + 1]
  const int k = blockIdx.x;
  if (k >= p.st->K) return;
  if (p.owned && !p.owned[k]) return;
  const int e0 = p.eptr[k], deg = p.eptr[k + 1] - e0;
  if (deg == 0) return;
  const int tid = threadIdx.x;
  const int Rw = 6 * deg, nb = (Rw + 1 + 7) / 8, Rp = nb * 8;
  const float *G = p.gram + p.goff[k];
  double *rhs = p.sys + (size_t)p.st->n_slots * GBA_NB2;
  const int pcr = p.kx[k] - p.t0;
  const int pc = (pcr >= 0 && pcr < p.P) ? p.perm[pcr] : -1;
  // position of each edge's destination pose in the permuted system: three dependent loads, once per edge
  int *s_pos = reinterpret_cast<int *>(s_U + 6 * (Rw + 1));
  for (int q = tid; q < deg; q += SCHUR_NT) {
    const int pj = p.geo[p.eidx[e0 + q]].pj;
    s_pos[q] = pj >= 0 ? p.perm[pj] : -1;
  }
  __syncthreads();
  auto pos_of = [&](int q) { return s_pos[q]; };

  // (a) edge-row x edge-row blocks, lower triangle only
  for (int idx = tid; idx < deg * deg; idx += SCHUR_NT) {
    const int n = idx / deg, m = idx - n * deg;
    const int a = pos_of(n), b = pos_of(m);
    if (a < 0 || b < 0 || a < b) continue;
    // all 36 values first (independent loads), then the slots of the <= 4 tiles the 6 x 6 block touches
    // (6 does not divide the tile size), then the adds -- not a dependent value + slot load per element
    float gv[36];
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c < 6; c++) gv[6 * r + c] = gram_at(G, Rp, 6 * n + r, 6 * m + c);
    const int R0 = 6 * a, C0 = 6 * b;
    const int tr0 = R0 / GBA_NB, tr1 = (R0 + 5) / GBA_NB, tc0 = C0 / GBA_NB, tc1 = (C0 + 5) / GBA_NB;
    const int s00 = p.tslot[tr0 * p.nT + tc0];
    const int s01 = (tc1 != tc0 && tr0 >= tc1) ? p.tslot[tr0 * p.nT + tc1] : s00;
    const int s10 = (tr1 != tr0) ? p.tslot[tr1 * p.nT + tc0] : s00;
    const int s11 = (tr1 != tr0 || tc1 != tc0) ? p.tslot[tr1 * p.nT + tc1] : s00;
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
      for (int c = 0; c < 6; c++) {
        const int Rg = R0 + r, Cg = C0 + c;
        if (Rg < Cg) continue;
        const bool hr = Rg / GBA_NB != tr0, hc = Cg / GBA_NB != tc0;
        const int slot = hr ? (hc ? s11 : s10) : (hc ? s01 : s00);
        atomicAdd(p.sys + (size_t)slot * GBA_NB2 + (Rg % GBA_NB) * GBA_NB + (Cg % GBA_NB), -(double)gv[6 * r + c]);
      }
  }
  // (b) rhs: v -= E Q w  (Ev6x1 + update_rhs, droid_kernels.cu:1298-1308)
  for (int idx = tid; idx < Rw; idx += SCHUR_NT) {
    const int n = idx / 6, r = idx - 6 * n;
    const int a = pos_of(n);
    if (a >= 0) atomicAdd(rhs + 6 * a + r, -(double)G[(size_t)idx * Rp + Rw]);
  }
  if (pc < 0) return;
  // (c) the keyframe's own pose row: U[:, col] = sum_n B_n G[n-block, col] for col in [0, Rw]
  for (int col = tid; col <= Rw; col += SCHUR_NT) {
    float u[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int n = 0; n < deg; n++) {
      const EdgeGeo *g = p.geo + p.eidx[e0 + n];
      float x[6], y[6];
      for (int r = 0; r < 6; r++) x[r] = gram_at(G, Rp, 6 * n + r, col);
      apply_B<float>(g->R, g->t, x, y);
      for (int r = 0; r < 6; r++) u[r] += y[r];
    }
    for (int r = 0; r < 6; r++) s_U[r * (Rw + 1) + col] = u[r];
  }
  __syncthreads();
  // cross blocks S[k, jj[m]] = U_m, stored on the lower side
  for (int idx = tid; idx < Rw; idx += SCHUR_NT) {
    const int m = idx / 6, c = idx - 6 * m;
    const int b = pos_of(m);
    if (b < 0 || b == pc) continue;
    for (int r = 0; r < 6; r++) {
      const double val = -(double)s_U[r * (Rw + 1) + idx];
      if (pc > b) atomicAdd(sys_at(p, 6 * pc + r, 6 * b + c), val);
      else atomicAdd(sys_at(p, 6 * b + c, 6 * pc + r), val);
    }
  }
  // diagonal block S[k,k] = sum_m U_m B_m^T and rhs
  if (tid < 6) {
    const int r = tid;
    float w[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int m = 0; m < deg; m++) {
      const EdgeGeo *g = p.geo + p.eidx[e0 + m];
      float x[6], y[6];
      for (int c = 0; c < 6; c++) x[c] = s_U[r * (Rw + 1) + 6 * m + c];
      apply_B<float>(g->R, g->t, x, y);
      for (int c = 0; c < 6; c++) w[c] += y[c];
    }
    for (int c = 0; c <= r; c++) atomicAdd(sys_at(p, 6 * pc + r, 6 * pc + c), -(double)w[c]);
    atomicAdd(rhs + 6 * pc + r, -(double)s_U[r * (Rw + 1) + Rw]);
  }
}

// Levenberg-Marquardt damping of the diagonal: d += ep + lm * d
// (droid_kernels.cu:1197 after the Schur complement; chol.py:66-68 before it).
// seg_n < 0: the whole system.  Edge-sharded (seg_n >= 1): the damping acts on the SUM over ranks, so by linearity
// this rank damps its own band segments [seg_lo, seg_lo + seg_n) in full (nobody else contributes to them), adds
// lm * d_local to the closure rows and -- on the first rank only -- their ep; foreign band rows are left alone.
__global__ void k_damp(BaParams p, float lm, float ep, int seg_lo, int seg_n, int first_rank, int nseg, int segT) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= 6 * p.P) return;
  const int i = 6 * p.perm[q / 6] + q % 6;
  double *d = sys_at(p, i, i);
  if (seg_n < 0) {
    *d += (double)ep + (double)lm * (*d);
    return;
  }
  const int tile = i / GBA_NB, band = nseg * segT;
  if (tile >= band) *d += (first_rank ? (double)ep : 0.0) + (double)lm * (*d);
  else if (tile / segT >= seg_lo && tile / segT < seg_lo + seg_n) *d += (double)ep + (double)lm * (*d);
}

// --------------------------------------------------------------------------------------------
// Pass C: back-substitution + disparity retraction, fused.  dz = Q (w - sum_rows E_r . dx_r):
// the E rows are rebuilt from target/weight (second streaming read) and contracted on the fly
// with one 6-vector per edge, delta_n = dx[jj] + B_n^T dx[k]; disps is updated in place.
// Replaces EvT6x1 + accum + disp_retr (droid_kernels.cu:1095-1115,1409-1418,933-946,1428-1431).
// --------------------------------------------------------------------------------------------
template <int NT, int NV, int MODE, bool IL>
__global__ void __launch_bounds__(NT, (NT >= 256) ? 3 : 6) k_backsub(BaParams p) {
  __shared__ EdgeGeo s_geo[LIN_EB];
  __shared__ float s_delta[LIN_EB][8];
  __shared__ int s_eid[LIN_EB];
  const int k = blockIdx.y;
  if (k >= p.st->K) return;
  if (p.owned && !p.owned[k]) return;
  const int chunk = blockIdx.x, tid = threadIdx.x;
  const int frame = p.kx[k], HW = p.HW;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const PixConst pk = make_pix_const(p.intr);
  const int pcr = frame - p.t0;
  // EvT6x1 skips pose index <= 0 (droid_kernels.cu:1105); chol.py:81 does not
  const int lo = (MODE == 0) ? 1 : 0;
  const bool kvalid = (pcr >= lo && pcr < p.P);

  // pixels are handled as PAIRS (px, px+1) with packed f32x2 arithmetic: pair q of group g = pixels 2q, 2q+1
  int px[NV];
  float y0[NV], z[NV][4];
  unsigned long long x2[NV][2], z2[NV][2], dw[NV][2];
#pragma unroll
  for (int g = 0; g < NV; g++) {
    px[g] = (chunk * NV + g) * (NT * 4) + tid * 4;
    float4 zz = make_float4(0.f, 0.f, 0.f, 0.f);
    if (px[g] < HW) zz = *reinterpret_cast<const float4 *>(p.disps + (size_t)frame * HW + px[g]);
    z[g][0] = zz.x; z[g][1] = zz.y; z[g][2] = zz.z; z[g][3] = zz.w;
    const float x0 = ((float)(px[g] % p.wd) - cx) / fx;
    y0[g] = ((float)(px[g] / p.wd) - cy) / fy;
#pragma unroll
    for (int q = 0; q < 2; q++) {
      x2[g][q] = pack2(fmaf((float)(2 * q), pk.ifx, x0), fmaf((float)(2 * q + 1), pk.ifx, x0));
      z2[g][q] = pack2(z[g][2 * q], z[g][2 * q + 1]);
      dw[g][q] = 0ull;
    }
  }
  const int e0 = p.eptr[k], e1 = p.eptr[k + 1];
  for (int eb = e0; eb < e1; eb += LIN_EB) {
    const int nb = min(LIN_EB, e1 - eb);
    __syncthreads();
    if (tid < nb) s_eid[tid] = p.eidx[eb + tid];
    __syncthreads();
    for (int i = tid; i < nb * 16; i += NT)
      reinterpret_cast<float *>(s_geo)[i] = reinterpret_cast<const float *>(p.geo + s_eid[i >> 4])[i & 15];
    __syncthreads();
    if (tid < nb) {
      const EdgeGeo &g = s_geo[tid];
      float d[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (g.pj >= lo)
        for (int r = 0; r < 6; r++) d[r] = p.dx[6 * g.pj + r];
      if (kvalid) {
        float y[6];
        apply_Bt(g.R, g.t, p.dx + 6 * pcr, y);
        for (int r = 0; r < 6; r++) d[r] += y[r];
      }
      if (g.flags & 1)
        for (int r = 0; r < 6; r++) d[r] = 0.f;
      for (int r = 0; r < 6; r++) s_delta[tid][r] = d[r];
    }
    __syncthreads();
    for (int q = 0; q < nb; q++) {
      const int n = s_eid[q];
      const float *dl = s_delta[q];
      const EdgeConst2 ec = make_edge_const2(s_geo[q].R, s_geo[q].t);
      // J . delta with the signs of ju' = (e, 0, -ae, -ab, opa, -b), jv' = (0, e, -be, -opb, ab, a) on the constants
      const unsigned long long D0 = dup2(dl[0]), D1 = dup2(dl[1]), nD2 = dup2(-dl[2]), nD3 = dup2(-dl[3]), D4 = dup2(dl[4]),
                               D5 = dup2(dl[5]), nD5 = dup2(-dl[5]);
      const float *wg = p.weights + (size_t)n * 2 * HW;
      float4 wu[NV], wv[NV];
#pragma unroll
      for (int g = 0; g < NV; g++) {
        if (px[g] < HW) ldg_uv4(wg, HW, px[g], IL, wu[g], wv[g]);  // weights only: E . delta carries no residual
        else wu[g] = wv[g] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int g = 0; g < NV; g++) {
        const unsigned long long wu2[2] = {pack2(wu[g].x, wu[g].y), pack2(wu[g].z, wu[g].w)};
        const unsigned long long wv2[2] = {pack2(wv[g].x, wv[g].y), pack2(wv[g].z, wv[g].w)};
#pragma unroll
        for (int h = 0; h < 2; h++) {
          PixHat2 L;
          pix_hat2<MODE>(ec, pk, x2[g][h], dup2(y0[g]), z2[g][h], wu2[h], wv2[h], L);
          const unsigned long long du =
              fma2r(L.e, D0, fma2r(L.ae, nD2, fma2r(L.ab, nD3, fma2r(L.opa, D4, mul2(L.b, nD5)))));
          const unsigned long long dv =
              fma2r(L.e, D1, fma2r(L.be, nD2, fma2r(L.opb, nD3, fma2r(L.ab, D4, mul2(L.a, D5)))));
          dw[g][h] = fma2r(mul2(L.wv, L.jzv), dv, fma2r(mul2(L.wu, L.jzu), du, dw[g][h]));
        }
      }
    }
  }
#pragma unroll
  for (int g = 0; g < NV; g++) {
    if (px[g] >= HW) continue;
    const float4 q4 = *reinterpret_cast<const float4 *>(p.Q + (size_t)k * HW + px[g]);
    const float4 w4 = *reinterpret_cast<const float4 *>(p.W + (size_t)k * HW + px[g]);
    const float qa[4] = {q4.x, q4.y, q4.z, q4.w}, wa[4] = {w4.x, w4.y, w4.z, w4.w};
    float dwa[4], dzv[4], zn[4];
    unpack2(dw[g][0], dwa[0], dwa[1]);
    unpack2(dw[g][1], dwa[2], dwa[3]);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      dzv[i] = qa[i] * (wa[i] - dwa[i]);
      zn[i] = z[g][i] + dzv[i];
      if (MODE == 1 && zn[i] < 0.f) zn[i] = 0.f;  // ba.py:119
    }
    *reinterpret_cast<float4 *>(p.disps + (size_t)frame * HW + px[g]) = make_float4(zn[0], zn[1], zn[2], zn[3]);
    if (p.dz_out)
      *reinterpret_cast<float4 *>(p.dz_out + (size_t)k * HW + px[g]) = make_float4(dzv[0], dzv[1], dzv[2], dzv[3]);
  }
}

// src/geom/ba.py:119 and :214 clamp the WHOLE disparity tensor at 0 after the update, also the keyframes
// the solve never touches (no out-edge): those rows (fmap < 0) are clamped here, the rows in kx by the
// update kernels themselves.  geom_py semantics and DSPO only; ba_cuda does not clamp.
__global__ void __launch_bounds__(128) k_clamp_rest(float *disps, const int *fmap, int HW) {
  const int frame = blockIdx.y;
  if (fmap[frame] >= 0) return;
  const int px = (blockIdx.x * 128 + threadIdx.x) * 4;
  if (px >= HW) return;
  float4 *q = reinterpret_cast<float4 *>(disps + (size_t)frame * HW + px);
  const float4 z = *q;
  if (z.x < 0.f || z.y < 0.f || z.z < 0.f || z.w < 0.f)
    *q = make_float4(fmaxf(z.x, 0.f), fmaxf(z.y, 0.f), fmaxf(z.z, 0.f), fmaxf(z.w, 0.f));
}

// SE3 retraction T <- Exp(xi) T for the optimised poses.
//   droid: droid_kernels.cu:110-175,877-931 (thresholds 1e-8 / 1e-4, no renormalisation)
//   geom : lietorch Exp + mul (so3.h:153-190, se3.h:134-142,45-47), quaternions renormalised
__global__ void k_pose_retr(BaParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.P) return;
  float *T = p.poses + 7 * (size_t)(p.t0 + i);
  const float *xi = p.dx + 6 * i;
  const float tau[3] = {xi[0], xi[1], xi[2]}, phi[3] = {xi[3], xi[4], xi[5]};
  const float th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float th = sqrtf(th2);
  float imag, real, dt[3] = {tau[0], tau[1], tau[2]};
  const bool small_q = (p.mode == 0) ? (th2 < 1e-8f) : (th < 1e-6f);
  if (small_q) {
    const float th4 = th2 * th2;
    imag = 0.5f - (1.0f / 48.0f) * th2 + (1.0f / 3840.0f) * th4;
    real = 1.0f - (1.0f / 8.0f) * th2 + (1.0f / 384.0f) * th4;
  } else {
    imag = sinf(0.5f * th) / th;
    real = cosf(0.5f * th);
  }
  float dq[4] = {imag * phi[0], imag * phi[1], imag * phi[2], real};
  float c1[3] = {phi[1] * tau[2] - phi[2] * tau[1], phi[2] * tau[0] - phi[0] * tau[2], phi[0] * tau[1] - phi[1] * tau[0]};
  float c2[3] = {phi[1] * c1[2] - phi[2] * c1[1], phi[2] * c1[0] - phi[0] * c1[2], phi[0] * c1[1] - phi[1] * c1[0]};
  float q[4] = {T[3], T[4], T[5], T[6]};
  if (p.mode == 0) {
    if (th > 1e-4f) {
      const float a = (1.f - cosf(th)) / th2, b = (th - sinf(th)) / (th * th2);
      for (int r = 0; r < 3; r++) dt[r] += a * c1[r] + b * c2[r];
    }
  } else {
    const float a = (th < 1e-6f) ? 0.5f - (1.0f / 24.0f) * th2 : (1.f - cosf(th)) / th2;
    const float b = (th < 1e-6f) ? (1.0f / 6.0f) - (1.0f / 120.0f) * th2 : (th - sinf(th)) / (th2 * th);
    for (int r = 0; r < 3; r++) dt[r] += a * c1[r] + b * c2[r];
    const float nd = rsqrtf(dq[0] * dq[0] + dq[1] * dq[1] + dq[2] * dq[2] + dq[3] * dq[3]);
    const float nq = rsqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int r = 0; r < 4; r++) {
      dq[r] *= nd;
      q[r] *= nq;
    }
  }
  float q1[4];
  q1[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  q1[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
  q1[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
  q1[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  if (p.mode == 1) {
    const float n1 = rsqrtf(q1[0] * q1[0] + q1[1] * q1[1] + q1[2] * q1[2] + q1[3] * q1[3]);
    for (int r = 0; r < 4; r++) q1[r] *= n1;
  }
  const float t[3] = {T[0], T[1], T[2]};
  const float uv0 = 2.f * (dq[1] * t[2] - dq[2] * t[1]);
  const float uv1 = 2.f * (dq[2] * t[0] - dq[0] * t[2]);
  const float uv2 = 2.f * (dq[0] * t[1] - dq[1] * t[0]);
  T[0] = t[0] + dq[3] * uv0 + (dq[1] * uv2 - dq[2] * uv1) + dt[0];
  T[1] = t[1] + dq[3] * uv1 + (dq[2] * uv0 - dq[0] * uv2) + dt[1];
  T[2] = t[2] + dq[3] * uv2 + (dq[0] * uv1 - dq[1] * uv0) + dt[2];
  T[3] = q1[0];
  T[4] = q1[1];
  T[5] = q1[2];
  T[6] = q1[3];
}
