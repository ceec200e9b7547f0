// gba_dmma.cuh -- fp64 tile kernels of the reduced-camera Cholesky on the TENSOR pipe (DMMA).
//
// The reference factorises the reduced camera system with Eigen::SimplicialLLT<double> on the host
// (droid_kernels.cu:1192-1213).  Here the 64 x 64 fp64 tile operations of the device-side tile Cholesky
// (gba_solver.cuh) -- triangular solve A X^T, symmetric update T -= L_i L_j^T and the factorisation of the
// diagonal tile itself -- are built from mma.sync.m8n8k4.f64 (SASS DMMA): one warp instruction is an
// 8 x 8 x 4 product, operands and accumulators in registers.  On B200 the fp64 tensor rate is about the
// rate of the fp64 CUDA-core pipe, but one DMMA replaces 8 DFMA plus their shared-memory operand loads, so a
// tile product stops being bound by shared-memory bandwidth and instruction issue.  Results are IEEE fp64
// fused multiply-adds in a different summation order.
//
// Shared-memory tiles use a pitch of 68 doubles: the 16 lanes of a half warp that load one DMMA fragment
// (4 rows x 4 consecutive k) then hit 16 different 8-byte banks.
#pragma once
#include "gba_common.cuh"

#define DM_LD 68
typedef double (*DmTile)[DM_LD];

// D (8x8) += A (8x4, row major fragment) * B (4x8, column major fragment).
// lane = 4 g + t:  a = A[g][t],  b = B[t][g],  d[0..1] = D[g][2t], D[g][2t+1]
__device__ __forceinline__ void dmma884(double (&d)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d[0]), "+d"(d[1])
               : "d"(a), "d"(b));
}

// Copy ROWS x 64 doubles from global (row stride 64) into a shared tile (pitch DM_LD); the 128-bit loads
// are all issued before the first store.
template <int ROWS>
__device__ __forceinline__ void dm_load_rows(DmTile dst, const double *src) {
  constexpr int NV = ROWS * GBA_NB / 2 / 256;
  static_assert(NV >= 1, "tile too small");
  double2 v[NV];
  const int tid = threadIdx.x;
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = __ldcg(reinterpret_cast<const double2 *>(src) + tid + 256 * k);
#pragma unroll
  for (int k = 0; k < NV; k++) {
    const int e = 2 * (tid + 256 * k);
    *reinterpret_cast<double2 *>(&dst[e >> 6][e & 63]) = v[k];
  }
}

__device__ __forceinline__ double dm_rcp(double d) {  // ~1e-12 relative: seed + one Newton step
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  const double e = fma(-d, r, 1.0);
  return fma(r, e, r);
}

// ---------------------------------------------------------------------------------------------------
// Factorisation of the 64 x 64 diagonal tile: X = L^-1 with A = L L^T, and y = X b.
// Block Gauss-Jordan on [A | I] with 8 x 8 blocks.  Per block column kb:
//   1. warp 0 (lanes 0-7 = the 8 rows, registers, shuffles): C = chol(D_kb)^-1 of the current diagonal
//      block -- the only dependent chain of the tile, 8 pivots of ~80 cycles;
//   2. all warps, DMMA: N = C [M rows of block kb]  (final rows of X),  G = A_panel C^T (= the panel of L);
//   3. DMMA: trailing A22 -= G G^T (lower blocks), M_below -= G N.  Warp 0 updates only the NEXT diagonal
//      block and goes straight on with its chain (look-ahead); warps 1-7 own one block row each and keep
//      up to 8 independent accumulators in flight (a single warp issues in order: a dependent DMMA costs
//      29 cycles, so the parallelism has to be spelled out).
// Measured on B200 (one CTA, tile in L2): 33.8 k cycles for the CUDA-core version it replaces.
// Shared memory: S (the tile, 64 x 68), M (the inverse being built, 64 x 68), G (64 x 8 panel), C (8 x 8).
// ---------------------------------------------------------------------------------------------------
struct DmPotrfSmem {
  double S[GBA_NB][DM_LD], M[GBA_NB][DM_LD], G[GBA_NB][12], C[8][12], b[GBA_NB];
  int bad;
};

#ifdef DM_TIMERS
__device__ long long dm_t[8];
__device__ long long dm_arr[8][8];
#define DM_T(i) { if (threadIdx.x == 0) { const long long now_ = clock64(); dm_t[i] += now_ - dm_last; dm_last = now_; } }
#else
#define DM_T(i)
#endif

// 1 / sqrt(d) to full precision without the library's slow path: fp32 seed + two Newton steps
__device__ __forceinline__ double dm_rsqrt(double d) {
  if (!(d > 1e-30 && d < 1e30)) return rsqrt(d);
  double y = (double)rsqrtf((float)d);
  y = y * fma(-0.5 * d, y * y, 1.5);
  y = y * fma(-0.5 * d, y * y, 1.5);
  return y;
}

// warp 0: C = chol(D)^-1 for the 8 x 8 block at (o, o) of S.  Lane r (mod 8) owns row r of D and of Lh^-1
// (D = Lh diag(d) Lh^T, Gauss-Jordan on [D | I]); rows travel by shuffle, all shuffles of a pivot are issued
// before the reciprocal they do not depend on.  ~1300 cycles.  Measured alternatives: two pivots per round
// (1960 cycles: more selects than it saves latency) and every lane factorising the whole block in its own
// registers without shuffles (1285 cycles, but 236 registers for the whole kernel).
__device__ __forceinline__ void dm_diag8(DmPotrfSmem &s, int o, int lane) {
  const int r = lane & 7;
  double a[8], m[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    a[k] = s.S[o + r][o + k];
    m[k] = 0.0;
  }
  double dmine = 1.0;
  bool bad = false;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    double aj[8], mj[8];
#pragma unroll
    for (int k = j; k < 8; k++) aj[k] = __shfl_sync(0xffffffffu, a[k], j);
#pragma unroll
    for (int c = 0; c < j; c++) mj[c] = __shfl_sync(0xffffffffu, m[c], j);
    const double dj = aj[j];
    if (!(dj > 0.0) || isinf(dj)) bad = true;
    if (r == j) dmine = dj;
    const double f = (r > j) ? a[j] * dm_rcp(dj) : 0.0;
#pragma unroll
    for (int k = j + 1; k < 8; k++) a[k] = fma(-f, aj[k], a[k]);
#pragma unroll
    for (int c = 0; c < j; c++) m[c] = fma(-f, mj[c], m[c]);
    m[j] = -f;  // M_j[j] = 1
  }
  if (__any_sync(0xffffffffu, bad)) {
    if (lane == 0) s.bad = 1;
  } else if (lane < 8) {
    const double sc = dm_rsqrt(dmine);
#pragma unroll
    for (int c = 0; c < 8; c++) s.C[r][c] = (c < r) ? sc * m[c] : (c == r ? sc : 0.0);
  }
}

// Lprev != nullptr: the tile left of this one, L[k,k-1], whose product has NOT been subtracted yet -- the update
// T - Lprev Lprev^T is applied here in shared memory (one hop less on the chain than a separate update task).
__device__ bool dm_potrf64(const double *T, double *XI, double *bvec, DmPotrfSmem &s, const double *Lprev = nullptr) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
#ifdef DM_TIMERS
  long long dm_last = clock64();
#endif
  if (tid == 0) s.bad = 0;
  if (tid < GBA_NB) s.b[tid] = __ldcg(bvec + tid);
  {  // tile -> S (coalesced 16-byte loads; only the lower triangle of T is valid), M = I
    double2 v[8], l[8];
#pragma unroll
    for (int q = 0; q < 8; q++) v[q] = __ldcg(reinterpret_cast<const double2 *>(T) + tid + 256 * q);
    if (Lprev) {  // both tiles in flight together: one L2 latency
#pragma unroll
      for (int q = 0; q < 8; q++) l[q] = __ldcg(reinterpret_cast<const double2 *>(Lprev) + tid + 256 * q);
    }
#pragma unroll
    for (int q = 0; q < 8; q++) {
      const int e = 2 * (tid + 256 * q), r = e >> 6, c = e & 63;
      s.S[r][c] = v[q].x;
      s.S[r][c + 1] = v[q].y;
      if (Lprev) {  // M holds L[k,k-1] for the moment
        s.M[r][c] = l[q].x;
        s.M[r][c + 1] = l[q].y;
      } else {
        s.M[r][c] = (r == c) ? 1.0 : 0.0;
        s.M[r][c + 1] = (r == c + 1) ? 1.0 : 0.0;
      }
    }
  }
  __syncthreads();
  if (Lprev) {
    // S -= M M^T on the lower block triangle (36 blocks of 8 x 8, K = 64): block rows r and 7 - r go to the warps
    // r and r + 4, 5 + 4 blocks, every block an independent DMMA chain
    const int r1 = warp & 3, r2 = 7 - r1, first = (warp < 4) ? 0 : 5, nblk = (warp < 4) ? 5 : 4;
    double acc[5][2];
#pragma unroll
    for (int q = 0; q < 5; q++) acc[q][0] = acc[q][1] = 0.0;
#pragma unroll 4
    for (int k = 0; k < GBA_NB; k += 4) {
#pragma unroll
      for (int q = 0; q < 5; q++) {
        if (q < nblk) {
          const int idx = first + q, bi = idx <= r1 ? r1 : r2, cj = idx <= r1 ? idx : idx - r1 - 1;
          dmma884(acc[q], s.M[8 * bi + g][k + t], s.M[8 * cj + g][k + t]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 5; q++) {
      if (q < nblk) {
        const int idx = first + q, bi = idx <= r1 ? r1 : r2, cj = idx <= r1 ? idx : idx - r1 - 1;
        double *dst = &s.S[8 * bi + g][8 * cj + 2 * t];
        dst[0] -= acc[q][0];
        dst[1] -= acc[q][1];
      }
    }
    __syncthreads();  // every warp is done reading M
#pragma unroll
    for (int q = 0; q < 8; q++) {
      const int e = 2 * (tid + 256 * q), r = e >> 6, c = e & 63;
      s.M[r][c] = (r == c) ? 1.0 : 0.0;
      s.M[r][c + 1] = (r == c + 1) ? 1.0 : 0.0;
    }
  }
  // the 8 x 8 diagonal blocks are read as full symmetric blocks: mirror their lower triangles
  for (int e = tid; e < 8 * 64; e += 256) {
    const int blk = e >> 6, r = (e >> 3) & 7, c = e & 7;
    if (c > r) s.S[8 * blk + r][8 * blk + c] = s.S[8 * blk + c][8 * blk + r];
  }
  __syncthreads();
  DM_T(0)
  if (warp == 0) dm_diag8(s, 0, lane);
  DM_T(7)
#pragma unroll 1
  for (int kb = 0; kb < 8; kb++) {
    const int o = 8 * kb, nb = 7 - kb;
#ifdef DM_TIMERS
    if (lane == 0) dm_arr[kb][warp] = clock64();
#endif
    __syncthreads();  // C of block kb is there, and every trailing update of step kb - 1 is done
    DM_T(1)
    if (s.bad) return false;
    // ---- 2. N = C M[o..o+8][0..o+8) (in place, final rows of X);  G = S[o+8..][o..o+8) C^T; one 8 x 8 block per job
    for (int job = warp; job < kb + 1 + nb; job += 8) {
      double acc[2] = {0.0, 0.0};
      if (job <= kb) {  // N block: rows o.., columns 8 job..: sum_k C[g][k] M[o + k][8 job + n]
        const double a0 = s.C[g][t], a1 = s.C[g][4 + t], b0 = s.M[o + t][8 * job + g], b1 = s.M[o + 4 + t][8 * job + g];
        dmma884(acc, a0, b0);
        dmma884(acc, a1, b1);
        __syncwarp();  // every lane has read its operands of M before the block is overwritten
        s.M[o + g][8 * job + 2 * t] = acc[0];
        s.M[o + g][8 * job + 2 * t + 1] = acc[1];
      } else {  // G block: rows o + 8 rb ..: sum_k S[row][o + k] C[n][k]
        const int rb = job - kb;
        const double a0 = s.S[o + 8 * rb + g][o + t], a1 = s.S[o + 8 * rb + g][o + 4 + t], b0 = s.C[g][t], b1 = s.C[g][4 + t];
        dmma884(acc, a0, b0);
        dmma884(acc, a1, b1);
        s.G[8 * rb + g][2 * t] = acc[0];
        s.G[8 * rb + g][2 * t + 1] = acc[1];
      }
    }
    __syncthreads();
    DM_T(2)
    if (nb == 0) break;
    // ---- 3. trailing update.  Block row bi (1..nb below the pivot block): S[bi][cj] -= G_bi G_cj^T for cj = 1..bi,
    //         M[bi][cj] -= G_bi N_cj for cj = 0..kb.
    if (warp == 0) {  // look-ahead: the next diagonal block only, then its factorisation
      double acc[2] = {0.0, 0.0};
      const double a0 = s.G[8 + g][t], a1 = s.G[8 + g][4 + t];
      dmma884(acc, a0, a0);  // B fragment of G_1 G_1^T is the same element
      dmma884(acc, a1, a1);
      double *dst = &s.S[o + 8 + g][o + 8 + 2 * t];
      dst[0] -= acc[0];
      dst[1] -= acc[1];
      __syncwarp();
      dm_diag8(s, o + 8, lane);
    }
    const int bi = warp;  // warps 1..7 own block rows 1..7 below the pivot block
    if (warp != 0 && bi <= nb) {
#ifdef DM_TIMERS
      long long w_t0 = clock64();
#endif
      const double a0 = s.G[8 * bi + g][t], a1 = s.G[8 * bi + g][4 + t];
      // S part: cj = 1..bi (block (1,1) is warp 0's); M part: cj = 0..kb.  All accumulators are independent.
      double accS[7][2], accM[8][2];
#pragma unroll
      for (int q = 0; q < 7; q++) {
        accS[q][0] = accS[q][1] = 0.0;
        const int cj = q + 1;
        if (cj <= bi && !(bi == 1 && cj == 1)) {
          dmma884(accS[q], a0, s.G[8 * cj + g][t]);
          dmma884(accS[q], a1, s.G[8 * cj + g][4 + t]);
        }
      }
#pragma unroll
      for (int cj = 0; cj < 8; cj++) {
        accM[cj][0] = accM[cj][1] = 0.0;
        if (cj <= kb) {
          dmma884(accM[cj], a0, s.M[o + t][8 * cj + g]);
          dmma884(accM[cj], a1, s.M[o + 4 + t][8 * cj + g]);
        }
      }
#ifdef DM_TIMERS
      long long w_t1 = clock64();
#endif
#pragma unroll
      for (int q = 0; q < 7; q++) {
        const int cj = q + 1;
        if (cj <= bi && !(bi == 1 && cj == 1)) {
          double *dst = &s.S[o + 8 * bi + g][o + 8 * cj + 2 * t];
          dst[0] -= accS[q][0];
          dst[1] -= accS[q][1];
        }
      }
#pragma unroll
      for (int cj = 0; cj < 8; cj++)
        if (cj <= kb) {
          double *dst = &s.M[o + 8 * bi + g][8 * cj + 2 * t];
          dst[0] -= accM[cj][0];
          dst[1] -= accM[cj][1];
        }
#ifdef DM_TIMERS
      if (tid == 32 * 4) { const long long w_t2 = clock64(); dm_t[5] += w_t1 - w_t0; dm_t[6] += w_t2 - w_t1; }
#endif
    }
    DM_T(3)
  }
  __syncthreads();
  // X = M (rows already scaled by the C blocks); y = X b
#pragma unroll
  for (int q = 0; q < 8; q++) {
    const int e = 2 * (tid + 256 * q), r = e >> 6, c = e & 63;
    double2 v;
    v.x = (c <= r) ? s.M[r][c] : 0.0;
    v.y = (c + 1 <= r) ? s.M[r][c + 1] : 0.0;
    reinterpret_cast<double2 *>(XI)[tid + 256 * q] = v;
  }
  {
    const int row = tid >> 2, part = tid & 3;
    double sa = 0.0, sb = 0.0;
#pragma unroll
    for (int k = part * 16; k < part * 16 + 16; k += 2) {
      sa = fma(s.M[row][k], s.b[k], sa);
      sb = fma(s.M[row][k + 1], s.b[k + 1], sb);
    }
    double sum = sa + sb;
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    if (part == 0) bvec[row] = sum;
  }
  DM_T(4)
  return true;
}

// ---------------------------------------------------------------------------------------------------
// Slab products of the tile Cholesky.  256 threads = 8 warps; warp w owns the 8 output columns 8w..8w+7
// of SLAB rows (RB = SLAB / 8 row blocks of 8), accumulators in registers.
// ---------------------------------------------------------------------------------------------------
#define DM_SMEM_BYTES (80 * 1024)  // >= DmPotrfSmem, >= two 64 x 68 tiles + the small vectors below

struct DmSlabSmem {
  double B[GBA_NB][DM_LD];  // the shared 64 x 64 operand: inv(L_kk) (trsm) or L[j,kk] (update)
  double A[GBA_NB][DM_LD];  // SLAB rows of the other operand
  double y[GBA_NB], part[8][GBA_NB];
};
static_assert(sizeof(DmSlabSmem) <= DM_SMEM_BYTES && sizeof(DmPotrfSmem) <= DM_SMEM_BYTES, "shared memory budget");

// acc[rb] += A[8 rb + g][k] * B[8 w + g'][k] over k in [0, kmax): the block row of A B^T this warp owns.
// Two interleaved accumulator sets when there are few row blocks, so that >= 4 DMMA chains are in flight.
template <int RB>
__device__ __forceinline__ void dm_slab_nt(double (&acc)[RB][2], const DmSlabSmem &s, int w, int kmax, int lane) {
  const int g = lane >> 2, t = lane & 3;
  if (RB <= 2) {
    double alt[RB][2];
#pragma unroll
    for (int rb = 0; rb < RB; rb++) alt[rb][0] = alt[rb][1] = 0.0;
#pragma unroll 2
    for (int k = 0; k < kmax; k += 8) {
      const double b0 = s.B[8 * w + g][k + t], b1 = s.B[8 * w + g][k + 4 + t];
#pragma unroll
      for (int rb = 0; rb < RB; rb++) {
        dmma884(acc[rb], s.A[8 * rb + g][k + t], b0);
        dmma884(alt[rb], s.A[8 * rb + g][k + 4 + t], b1);
      }
    }
#pragma unroll
    for (int rb = 0; rb < RB; rb++) {
      acc[rb][0] += alt[rb][0];
      acc[rb][1] += alt[rb][1];
    }
  } else {
#pragma unroll 2
    for (int k = 0; k < kmax; k += 4) {
      const double b = s.B[8 * w + g][k + t];
#pragma unroll
      for (int rb = 0; rb < RB; rb++) dmma884(acc[rb], s.A[8 * rb + g][k + t], b);
    }
  }
}
