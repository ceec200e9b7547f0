// gba_dspo.cuh -- DSPO stage 2 (disparity + per-keyframe scale/shift of the mono prior).
//
// Reference: src/geom/ba.py:127-216 (BA_with_scale_shift) + src/geom/chol.py:58-85.  The
// reference scatters H_wq / E_wq_d with (ll, ll) into dense [M,M,2,2] / [M,M,2,HW] tensors
// (ba.py:191-192,203-204), i.e. the system is block diagonal: every keyframe is an independent
// 2x2 Schur system.  Here it is one streaming pass over the edges (J_z terms only), a per-frame
// reduction of 7 sums, M tiny 2x2 solves and one per-pixel update pass; nothing of size M^2.
//
// Numerics: S = H - E Q E^T cancels catastrophically in fp32 when the prior dominates
// (J_d^2 Q -> 1).  The kernels accumulate the algebraically identical, cancellation-free form
//   H_ab - (E Q E^T)_ab = sum_p J_a J_b (C_proj + eta) Q,
//   u_a - (E Q w)_a     = -sum_p J_a Q (r_d (C_proj + eta) + J_d w_reproj).
#pragma once
#include "gba_common.cuh"

struct DspoParams {
  const float *poses;
  float *disps;
  const float *intr, *targets, *weights, *eta, *mono;
  float *scales, *shifts;
  const unsigned char *mask;
  const int64_t *ii, *jj;
  int N, E, HW, wd, ignore_frames, interleaved, nchunk;
  float lm, ep, sqrt_alpha;
  int *kx, *eptr, *eidx;
  Status *st;
  EdgeGeo *geo;
  float *Q, *W;
  float *part;  // [K][nchunk][8]
  float *dwq;   // [K][2]
  int *bad;
};

template <bool IL>
__device__ __forceinline__ void load_tw4(const float *tg, const float *wg, int n, int HW, int px, float (&tu)[4],
                                         float (&tv)[4], float (&wu)[4], float (&wv)[4]) {
  if (IL) {  // [E,ht,wd,2]
    const float4 a = ldg_stream4(tg + ((size_t)n * HW + px) * 2), b = ldg_stream4(tg + ((size_t)n * HW + px) * 2 + 4);
    const float4 c = ldg_stream4(wg + ((size_t)n * HW + px) * 2), d = ldg_stream4(wg + ((size_t)n * HW + px) * 2 + 4);
    tu[0] = a.x; tv[0] = a.y; tu[1] = a.z; tv[1] = a.w; tu[2] = b.x; tv[2] = b.y; tu[3] = b.z; tv[3] = b.w;
    wu[0] = c.x; wv[0] = c.y; wu[1] = c.z; wv[1] = c.w; wu[2] = d.x; wv[2] = d.y; wu[3] = d.z; wv[3] = d.w;
  } else {  // [E,2,ht,wd]
    const float4 a = ldg_stream4(tg + (size_t)n * 2 * HW + px), b = ldg_stream4(tg + (size_t)n * 2 * HW + HW + px);
    const float4 c = ldg_stream4(wg + (size_t)n * 2 * HW + px), d = ldg_stream4(wg + (size_t)n * 2 * HW + HW + px);
    tu[0] = a.x; tu[1] = a.y; tu[2] = a.z; tu[3] = a.w; tv[0] = b.x; tv[1] = b.y; tv[2] = b.z; tv[3] = b.w;
    wu[0] = c.x; wu[1] = c.y; wu[2] = c.z; wu[3] = c.w; wv[0] = d.x; wv[1] = d.y; wv[2] = d.z; wv[3] = d.w;
  }
}

// prior Jacobians of one pixel (ba.py:145-168)
__device__ __forceinline__ void dspo_prior(float mono, bool vdm, bool ignore, float sa, float &Jd, float &Js,
                                           float &Jq) {
  const bool inval = (mono < 1e-6f) || ignore;
  const float sap = vdm ? sa * 10.f : sa;
  Jd = (inval && vdm) ? 0.f : sap;
  Js = inval ? 0.f : -mono * sap;
  Jq = inval ? 0.f : -sap;
}

template <int NT, bool IL>
__global__ void __launch_bounds__(NT) k_dspo_lin(DspoParams p) {
  constexpr int NW = NT / 32;
  __shared__ float s_red[NW][8];
  const int k = blockIdx.y;
  if (k >= p.st->K) return;
  const int chunk = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int frame = p.kx[k], HW = p.HW;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const int px = chunk * (NT * 4) + tid * 4;
  float sums[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (px < HW) {
    float x0[4], y0[4], z[4], c[4] = {0.f, 0.f, 0.f, 0.f}, wr[4] = {0.f, 0.f, 0.f, 0.f};
    const float4 zz = *reinterpret_cast<const float4 *>(p.disps + (size_t)frame * HW + px);
    z[0] = zz.x; z[1] = zz.y; z[2] = zz.z; z[3] = zz.w;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      x0[i] = ((float)((px + i) % p.wd) - cx) / fx;
      y0[i] = ((float)((px + i) / p.wd) - cy) / fy;
    }
    for (int e = p.eptr[k]; e < p.eptr[k + 1]; e++) {
      const int n = p.eidx[e];
      const EdgeGeo *g = p.geo + n;
      float R[9], t[3];
      {
        const float4 g0 = __ldg(reinterpret_cast<const float4 *>(g)), g1 = __ldg(reinterpret_cast<const float4 *>(g) + 1),
                     g2 = __ldg(reinterpret_cast<const float4 *>(g) + 2);
        R[0] = g0.x; R[1] = g0.y; R[2] = g0.z; R[3] = g0.w;
        R[4] = g1.x; R[5] = g1.y; R[6] = g1.z; R[7] = g1.w;
        R[8] = g2.x; t[0] = g2.y; t[1] = g2.z; t[2] = g2.w;
      }
      float tu[4], tv[4], wu[4], wv[4];
      load_tw4<IL>(p.targets, p.weights, n, HW, px, tu, tv, wu, wv);
#pragma unroll
      for (int i = 0; i < 4; i++) {
        PixLin L;
        pix_linearize(R, t, fx, fy, cx, cy, 1, x0[i], y0[i], z[i], tu[i], tv[i], wu[i], wv[i], L);
        c[i] += L.wu * L.jzu * L.jzu + L.wv * L.jzv * L.jzv;  // Ck, ba.py:186
        wr[i] += L.wu * L.ru * L.jzu + L.wv * L.rv * L.jzv;   // -wk, ba.py:185
      }
    }
    const float4 mo4 = *reinterpret_cast<const float4 *>(p.mono + (size_t)frame * HW + px);
    const float4 et4 = *reinterpret_cast<const float4 *>(p.eta + (size_t)k * HW + px);
    const uchar4 mk4 = *reinterpret_cast<const uchar4 *>(p.mask + (size_t)frame * HW + px);
    const float mo[4] = {mo4.x, mo4.y, mo4.z, mo4.w}, et[4] = {et4.x, et4.y, et4.z, et4.w};
    const bool mk[4] = {mk4.x != 0, mk4.y != 0, mk4.z != 0, mk4.w != 0};
    const float s = p.scales[frame], q = p.shifts[frame];
    const bool ignore = frame < p.ignore_frames;
    float qo[4], wo[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float Jd, Js, Jq;
      dspo_prior(mo[i], mk[i], ignore, p.sqrt_alpha, Jd, Js, Jq);
      const float rdep = p.sqrt_alpha * (z[i] - (s * mo[i] + q));  // un-boosted sqrt(alpha), ba.py:155
      const float cp = c[i] + et[i];
      const float C = cp + Jd * Jd;       // ba.py:199
      const float ww = wr[i] - Jd * rdep;  // ba.py:202
      const float Q = 1.0f / C;
      qo[i] = Q;
      wo[i] = ww;
      const float cq = cp * Q;
      sums[0] += Js * Js * cq;
      sums[1] += Js * Jq * cq;
      sums[2] += Jq * Jq * cq;
      sums[3] += Js * Js;
      sums[4] += Jq * Jq;
      const float rr = Q * (rdep * cp + Jd * wr[i]);
      sums[5] -= Js * rr;
      sums[6] -= Jq * rr;
    }
    *reinterpret_cast<float4 *>(p.Q + (size_t)k * HW + px) = make_float4(qo[0], qo[1], qo[2], qo[3]);
    *reinterpret_cast<float4 *>(p.W + (size_t)k * HW + px) = make_float4(wo[0], wo[1], wo[2], wo[3]);
  }
#pragma unroll
  for (int v = 0; v < 7; v++) {
    float x = sums[v];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) s_red[warp][v] = x;
  }
  __syncthreads();
  if (tid < 7) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < NW; w++) x += s_red[w][tid];
    p.part[((size_t)k * p.nchunk + chunk) * 8 + tid] = x;
  }
}

// per-keyframe 2x2 solve; one non-SPD block zeroes every (scale, shift) update, as the single
// dense Cholesky of the reference does when it throws (chol.py:10-15)
__global__ void __launch_bounds__(256) k_dspo_solve(DspoParams p) {
  __shared__ int s_bad;
  const int K = p.st->K;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    double s[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < p.nchunk; c++)
      for (int v = 0; v < 7; v++) s[v] += (double)p.part[((size_t)k * p.nchunk + c) * 8 + v];
    const double a00 = s[0] + (double)p.ep + (double)p.lm * s[3];  // damping on H before Schur, chol.py:66-68
    const double a11 = s[2] + (double)p.ep + (double)p.lm * s[4];
    const double a01 = s[1];
    bool bad = !(a00 > 0.0);
    const double l00 = bad ? 1.0 : sqrt(a00);
    const double l10 = a01 / l00;
    if (!(a11 - l10 * l10 > 0.0)) bad = true;
    if (bad) {
      s_bad = 1;
      continue;
    }
    const double l11 = sqrt(a11 - l10 * l10);
    const double y0 = s[5] / l00, y1 = (s[6] - l10 * y0) / l11;
    const double x1 = y1 / l11, x0 = (y0 - l10 * x1) / l00;
    p.dwq[2 * k] = (float)x0;
    p.dwq[2 * k + 1] = (float)x1;
  }
  __syncthreads();
  if (s_bad) {
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
      p.dwq[2 * k] = 0.f;
      p.dwq[2 * k + 1] = 0.f;
    }
    if (threadIdx.x == 0) {
      *p.bad = 1;
      p.st->chol_fail += 1;
    }
  }
}

// dz = Q (w - E^T dwq); disps <- clamp(disps + dz, 0); (s, q) <- (s, q) + dwq   (ba.py:206-214)
template <int NT>
__global__ void __launch_bounds__(NT) k_dspo_update(DspoParams p) {
  const int k = blockIdx.y;
  if (k >= p.st->K) return;
  const int frame = p.kx[k], HW = p.HW;
  const int px = blockIdx.x * (NT * 4) + threadIdx.x * 4;
  const float ds = p.dwq[2 * k], dq = p.dwq[2 * k + 1];
  if (px < HW) {
    const float4 mo4 = *reinterpret_cast<const float4 *>(p.mono + (size_t)frame * HW + px);
    const uchar4 mk4 = *reinterpret_cast<const uchar4 *>(p.mask + (size_t)frame * HW + px);
    const float4 q4 = *reinterpret_cast<const float4 *>(p.Q + (size_t)k * HW + px);
    const float4 w4 = *reinterpret_cast<const float4 *>(p.W + (size_t)k * HW + px);
    float4 z4 = *reinterpret_cast<const float4 *>(p.disps + (size_t)frame * HW + px);
    const float mo[4] = {mo4.x, mo4.y, mo4.z, mo4.w}, qa[4] = {q4.x, q4.y, q4.z, q4.w}, wa[4] = {w4.x, w4.y, w4.z, w4.w};
    const bool mk[4] = {mk4.x != 0, mk4.y != 0, mk4.z != 0, mk4.w != 0};
    float z[4] = {z4.x, z4.y, z4.z, z4.w};
    const bool ignore = frame < p.ignore_frames;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float Jd, Js, Jq;
      dspo_prior(mo[i], mk[i], ignore, p.sqrt_alpha, Jd, Js, Jq);
      const float dz = qa[i] * (wa[i] - (Js * Jd * ds + Jq * Jd * dq));
      z[i] = fmaxf(z[i] + dz, 0.f);
    }
    *reinterpret_cast<float4 *>(p.disps + (size_t)frame * HW + px) = make_float4(z[0], z[1], z[2], z[3]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    p.scales[frame] += ds;
    p.shifts[frame] += dq;
  }
}

