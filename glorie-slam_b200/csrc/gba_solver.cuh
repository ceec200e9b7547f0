// gba_solver.cuh -- reduced camera system: ordering, tile structure, fp64 tile Cholesky, solve.
//
// The reference factorises the 6P x 6P system on the HOST: D2H copy, fp32->fp64, triplet
// assembly and Eigen::SimplicialLLT<double> with AMD ordering every iteration
// (droid_kernels.cu:1131-1213).  Here everything stays on the device:
//   * an "arrowhead" pose ordering (temporal band first, loop-closure poses last) plays the role
//     of the fill-reducing ordering;
//   * the matrix is stored as fp64 NB x NB tiles of the lower triangle, only the tiles that are
//     structurally non-zero after a bit-parallel symbolic elimination exist;
//   * one fused kernel per tile column does Cholesky of the diagonal tile, the triangular solves
//     of the tiles below it, the trailing update and the forward substitution of the rhs, the
//     three phases chained by in-kernel flags instead of kernel boundaries;
//   * the backward substitution is one small launch per tile row.
// The numbers are those of a dense fp64 Cholesky of the permuted matrix.
#pragma once
#include "gba_common.cuh"

#define TS_MAXT 512  // max tile rows (6 * P <= 32768)
#define TS_WORDS (TS_MAXT / 32)

struct SolverBufs {
  unsigned int *tbits;    // [nT][TS_WORDS] structure bits, row i bit j (j <= i)
  int *colrows, *colcnt;  // active tiles below the diagonal, per tile column
  int *rowcols, *rowcnt;  // active tiles left of the diagonal, per tile row
  double *invL;           // [nT][NB*NB]
  int *fail;              // factorisation hit a non-positive pivot
  unsigned int *flagA, *cntB, *cntD, *flagX, *ticket;  // per-panel in-kernel synchronisation
  unsigned int *upd;      // [n_slots] slab updates a tile has received inside k_chain
};

// ---------------------------------------------------------------------------------------------
// Pose ordering = arrowhead + nested dissection, chosen on the device (one CTA).
//  1. Edges with |i - j| > W are "long range"; their destination poses form the closure set C
//     (through the Schur complement the destination couples with every neighbour of the source, so
//     it cannot stay inside a band).  W is picked from a few candidates by a flop model of the
//     resulting band + arrow factorisation.
//  2. The remaining poses form a band (temporal order) of half width hb = 2 * max short-range
//     |i - j|.  The band is cut into n_seg segments by separators of hb consecutive band poses; the
//     separators join the closure block.  Segments are independent of each other, so the same
//     step of every segment is factorised by one launch: the sequential chain shrinks from
//     all tile columns to seg_tiles + closure tile columns.
//  3. Segments start on a multiple of 32 pose positions (= 3 tiles exactly) so that no tile is
//     shared by two segments; the unused positions are padding (unit diagonal).
// Layout of positions: [segment 0 | pad][segment 1 | pad] ... [separators and C, temporal order].
// ---------------------------------------------------------------------------------------------
#define ORD_NC 8
#define ORD_ALIGN 32  // pose positions; 32 * 6 rows = 3 tiles
__global__ void __launch_bounds__(1024) k_order(BaParams p, int pos_cap) {
  __shared__ int s_cnt[ORD_NC], s_reff[ORD_NC];
  __shared__ int s_best, s_warp[32], s_nseg, s_seglen, s_hb, s_SL;
  extern __shared__ unsigned int s_flagw[];  // one byte per pose: bit c = in C for candidate c
  unsigned char *s_flag = reinterpret_cast<unsigned char *>(s_flagw);
  const int tid = threadIdx.x, P = p.P;
  const int cand[ORD_NC] = {4, 8, 16, 32, 64, 128, 256, 1 << 30};
  for (int i = tid; i < (P + 3) / 4; i += blockDim.x) s_flagw[i] = 0;
  if (tid < ORD_NC) {
    s_cnt[tid] = 0;
    s_reff[tid] = 0;
  }
  __syncthreads();
  for (int n = tid; n < p.E; n += blockDim.x) {
    const long long i = p.ii[n], j = p.jj[n];
    const long long pj = j - p.t0;
    if (pj < 0 || pj >= P || i == j) continue;
    const int d = (int)(i > j ? i - j : j - i);
    unsigned m = 0;
#pragma unroll
    for (int c = 0; c < ORD_NC; c++) {
      if (d > cand[c]) m |= 1u << c;
      else atomicMax(&s_reff[c], d);
    }
    if (m) atomicOr(&s_flagw[pj >> 2], m << (8 * (pj & 3)));
  }
  __syncthreads();
  for (int i = tid; i < P; i += blockDim.x) {
    const unsigned m = s_flag[i];
#pragma unroll
    for (int c = 0; c < ORD_NC; c++)
      if (m & (1u << c)) atomicAdd(&s_cnt[c], 1);
  }
  __syncthreads();
  if (tid == 0) {
    double best = 1e300;
    int bi = ORD_NC - 1;
    for (int c = 0; c < ORD_NC; c++) {
      const double nC = s_cnt[c], nB = P - nC;
      const double b = fmin(2.0 * (double)max(s_reff[c], 1), nB);  // band half width in poses
      // band factor + arrow rows + arrow trailing update + dense tail (pose^3 units)
      const double cost = nB * b * b + nC * nB * b + nB * nC * nC * 0.5 + nC * nC * nC / 3.0;
      if (cost < best) {
        best = cost;
        bi = c;
      }
    }
    s_best = bi;
    const int nC = s_cnt[bi], nB = P - nC, hb = min(2 * max(s_reff[bi], 1), nB);
    // nested dissection: pick the segment count with the shortest sequential chain of tile columns
    int best_seg = 1, best_chain = 1 << 30, best_len = nB;
    for (int ns = 1; ns <= 16; ns *= 2) {
      const int nsep = (ns - 1) * hb;
      if (ns > 1 && (nB - nsep < ns * 2 * hb || nB - nsep < ns * ORD_ALIGN)) break;  // segments too short
      const int len = (nB - nsep + ns - 1) / ns;                               // real poses per segment
      const int SL = (ns == 1) ? len : (len + ORD_ALIGN - 1) / ORD_ALIGN * ORD_ALIGN;
      if ((ns == 1 ? nB : ns * SL) + nC + nsep > pos_cap) break;
      const int chain = (6 * SL + GBA_NB - 1) / GBA_NB + (6 * (nC + nsep) + GBA_NB - 1) / GBA_NB;
      if (chain < best_chain) {
        best_chain = chain;
        best_seg = ns;
        best_len = len;
      }
    }
    s_nseg = best_seg;
    s_seglen = best_len;
    s_hb = hb;
    s_SL = (best_seg == 1) ? best_len : (best_len + ORD_ALIGN - 1) / ORD_ALIGN * ORD_ALIGN;
  }
  __syncthreads();
  const int bc = s_best, nseg = s_nseg, seglen = s_seglen, hb = s_hb, SL = s_SL;
  const int band_end = nseg * SL;  // first position of the closure block (nseg == 1: == nB)
  // band index b of a non-C pose = number of non-C poses before it.  With nseg > 1 the band is
  // [seg 0: seglen][sep: hb][seg 1: seglen][sep: hb] ... ; separators go to the closure block.
  // Closure positions are handed out in temporal order: a second scan counts closure members.
  for (int i = tid; i < pos_cap; i += blockDim.x) p.iperm[i] = -1;
  __syncthreads();
  int carry_band = 0, carry_clo = 0;
  for (int base = 0; base < P; base += blockDim.x) {
    const int i = base + tid;
    const int inC = (i < P) ? ((s_flag[i] >> bc) & 1) : 0;
    // pass 1: band index
    int v = (i < P) ? 1 - inC : 0;
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
    const int bincl = v + ((tid >> 5) ? s_warp[(tid >> 5) - 1] : 0) + carry_band;
    const int total_band = s_warp[31] + carry_band;
    __syncthreads();
    // is this pose a closure member (C or separator)?
    int seg = 0, off = 0, clo = inC;
    if (i < P && !inC) {
      const int bidx = bincl - 1;
      if (nseg > 1) {
        seg = bidx / (seglen + hb);
        off = bidx - seg * (seglen + hb);
        if (seg >= nseg) {  // tail beyond the last separator belongs to the last segment
          seg = nseg - 1;
          off = bidx - seg * (seglen + hb);
        }
        if (off >= seglen && seg < nseg - 1) clo = 1;      // separator
        else if (off >= SL) clo = 1;                        // overflow of the last segment
      } else
        off = bidx;
    }
    // pass 2: closure index
    int w2 = (i < P) ? clo : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, w2, o);
      if ((tid & 31) >= o) w2 += u;
    }
    if ((tid & 31) == 31) s_warp[tid >> 5] = w2;
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
    const int cincl = w2 + ((tid >> 5) ? s_warp[(tid >> 5) - 1] : 0) + carry_clo;
    const int total_clo = s_warp[31] + carry_clo;
    __syncthreads();
    if (i < P) {
      const int pos = clo ? band_end + cincl - 1 : seg * SL + off;
      p.perm[i] = pos;
      p.iperm[pos] = i;
    }
    carry_band = total_band;
    carry_clo = total_clo;
  }
  if (tid == 0) {
    const int npos = band_end + carry_clo;
    p.st->n_closure = carry_clo;
    p.st->band_w = cand[bc] > 1000000 ? -1 : cand[bc];
    p.st->n_seg = nseg;
    p.st->seg_tiles = (nseg == 1) ? 0 : 6 * SL / GBA_NB;
    p.st->n_pos = npos;
    p.st->n_tiles = (6 * npos + GBA_NB - 1) / GBA_NB;
  }
}

// Mark the tiles touched by the pose pairs of one depth keyframe (its own pose and the
// destination poses of its out-edges are all mutually coupled through the Schur complement;
// the pose-pose Hessian blocks (ii,ii),(ii,jj),(jj,jj) are a subset).
template <bool CHECK>
__global__ void k_tile_mark(BaParams p, unsigned int *tbits) {
  const int k = blockIdx.x;
  if (k >= p.st->K) return;
  const int e0 = p.eptr[k], deg = p.eptr[k + 1] - e0;
  const int L = deg + 1;
  auto pose_of = [&](int q) {
    const int pr = (q < deg) ? (int)p.jj[p.eidx[e0 + q]] - p.t0 : p.kx[k] - p.t0;
    return (pr >= 0 && pr < p.P) ? p.perm[pr] : -1;
  };
  for (int idx = threadIdx.x; idx < L * L; idx += blockDim.x) {
    const int a = pose_of(idx / L), b = pose_of(idx % L);
    if (a < 0 || b < 0 || a < b) continue;
    const int r0 = (6 * a) / GBA_NB, r1 = (6 * a + 5) / GBA_NB;
    const int c0 = (6 * b) / GBA_NB, c1 = (6 * b + 5) / GBA_NB;
    for (int r = r0; r <= r1; r++)
      for (int c = c0; c <= c1; c++)
        if (r >= c) {
          if (!CHECK) atomicOr(&tbits[r * TS_WORDS + (c >> 5)], 1u << (c & 31));
          else if (!((tbits[r * TS_WORDS + (c >> 5)] >> (c & 31)) & 1u)) atomicOr(&p.st->err, 8);
        }
  }
}

// One CTA: symbolic fill on bit rows (kept in global memory, L1/L2 resident), slot numbering,
// column / row lists.
__global__ void __launch_bounds__(512) k_tile_fill(BaParams p, SolverBufs sb) {
  __shared__ unsigned int s_col[TS_WORDS];
  __shared__ int s_start[TS_MAXT + 1];
  __shared__ int s_maxc, s_maxr;
  unsigned int *bits = sb.tbits;
  const int nT = p.st->n_tiles, nS = p.nT, tid = threadIdx.x, nW = (nT + 31) / 32;  // nS = table stride
  for (int i = tid; i < nT; i += blockDim.x) bits[i * TS_WORDS + (i >> 5)] |= 1u << (i & 31);
  if (tid == 0) s_maxc = s_maxr = 0;
  __syncthreads();
  // right-looking elimination: rows i >= j in column kk's structure create tile (i, j)
  for (int kk = 0; kk < nT; kk++) {
    for (int w = tid; w < nW; w += blockDim.x) s_col[w] = 0;
    __syncthreads();
    for (int i = kk + 1 + tid; i < nT; i += blockDim.x)
      if ((bits[i * TS_WORDS + (kk >> 5)] >> (kk & 31)) & 1) atomicOr(&s_col[i >> 5], 1u << (i & 31));
    __syncthreads();
    for (int i = kk + 1 + tid; i < nT; i += blockDim.x) {
      if (!((s_col[i >> 5] >> (i & 31)) & 1)) continue;
      for (int w = 0; w <= (i >> 5); w++) {
        unsigned m = s_col[w];
        if (w == (i >> 5)) m &= ((i & 31) == 31) ? 0xffffffffu : ((2u << (i & 31)) - 1u);
        bits[i * TS_WORDS + w] |= m;
      }
    }
    __syncthreads();
  }
  // slots: row-major over the lower triangle
  for (int i = tid; i < nT; i += blockDim.x) {
    int c = 0;
    for (int w = 0; w < nW; w++) c += __popc(bits[i * TS_WORDS + w]);
    s_start[i + 1] = c;
  }
  __syncthreads();
  if (tid == 0) {
    s_start[0] = 0;
    for (int i = 0; i < nT; i++) s_start[i + 1] += s_start[i];
    p.st->n_slots = s_start[nT];
  }
  __syncthreads();
  for (int idx = tid; idx < nT * nT; idx += blockDim.x) {
    const int i = idx / nT, j = idx % nT;
    int slot = -1;
    if (j <= i && ((bits[i * TS_WORDS + (j >> 5)] >> (j & 31)) & 1)) {
      int c = 0;
      for (int w = 0; w < (j >> 5); w++) c += __popc(bits[i * TS_WORDS + w]);
      c += __popc(bits[i * TS_WORDS + (j >> 5)] & ((1u << (j & 31)) - 1u));
      slot = s_start[i] + c;
    }
    p.tslot[i * nS + j] = slot;
  }
  for (int kk = tid; kk < nT; kk += blockDim.x) {
    int c = 0;
    for (int i = kk + 1; i < nT; i++)
      if ((bits[i * TS_WORDS + (kk >> 5)] >> (kk & 31)) & 1) sb.colrows[kk * nS + c++] = i;
    sb.colcnt[kk] = c;
    atomicMax(&s_maxc, c);
    int r = 0;
    for (int j = 0; j < kk; j++)
      if ((bits[kk * TS_WORDS + (j >> 5)] >> (j & 31)) & 1) sb.rowcols[kk * nS + r++] = j;
    sb.rowcnt[kk] = r;
    atomicMax(&s_maxr, r);
  }
  __syncthreads();
  if (tid == 0) {
    p.st->max_col = s_maxc;
    p.st->max_row = s_maxr;
  }
}

// rows of padding positions (segment alignment, tail of the last tile) get a unit diagonal so the
// factorisation runs through them
__global__ void k_pad_diag(BaParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.st->n_tiles * GBA_NB) return;
  const int pos = i / 6;
  if (pos >= p.st->n_pos || p.iperm[pos] < 0) *sys_at(p, i, i) = 1.0;
}

// ---------------------------------------------------------------------------------------------
// Tile-column roles (256 threads; dynamic shared memory: 2 tiles of 64 x 65 doubles + 16 rows).
// fp64 runs at 64 FMA/clk/SM on B200, so a 64^3 tile product pins one SM for ~9 us; the
// trsm / syrk roles therefore work on SLAB-row slabs of a tile (SPLIT = 64 / SLAB CTAs per tile)
// to spread a tile column over more SMs, and the potrf role keeps divisions and square roots off
// its critical path (Newton rsqrt from an fp32 seed, blocked inverse without divisions).
// ---------------------------------------------------------------------------------------------
typedef double (*TilePtr)[GBA_NB + 1];
#define TILE_STRIDE (GBA_NB * (GBA_NB + 1))

__device__ __forceinline__ double *rhs_ptr(const BaParams &p) { return p.sys + (size_t)p.st->n_slots * GBA_NB2; }

__device__ __forceinline__ double fast_rsqrt(double d) {
  double y = (double)rsqrtf((float)d);
  y = y * (1.5 - 0.5 * d * y * y);
  y = y * (1.5 - 0.5 * d * y * y);
  return y;
}

// Copy ROWS x 64 doubles from global (L2, row stride 64) into a padded shared tile; all the
// 128-bit loads are issued before the first store so that their latencies overlap.
template <int ROWS>
__device__ __forceinline__ void load_rows(TilePtr dst, const double *src) {
  constexpr int NV = ROWS * GBA_NB / 2 / 256;  // double2 per thread
  static_assert(NV >= 1, "tile too small");
  double2 v[NV];
  const int tid = threadIdx.x;
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = __ldcg(reinterpret_cast<const double2 *>(src) + tid + 256 * k);
#pragma unroll
  for (int k = 0; k < NV; k++) {
    const int e = 2 * (tid + 256 * k);
    dst[e >> 6][e & 63] = v[k].x;
    dst[e >> 6][(e & 63) + 1] = v[k].y;
  }
}

// Reciprocal for the pivot loop.  Dependent fp64 operations cost ~40 cycles each on B200 (measured
// with clock64), so the chain is kept to seed + ONE Newton step: rcp.approx.ftz.f64 gives ~20 bits,
// one step ~40 bits (1e-12 relative).  The trailing update therefore eliminates with a pivot that is
// off by 1e-12 relative -- a backward error far below fp64 assembly noise; the columns of L are
// scaled afterwards with an exact sqrt of the pivots actually used.
__device__ __forceinline__ double fast_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  const double e = fma(-d, r, 1.0);
  return fma(r, e, r);
}

// Cholesky of the diagonal tile, its inverse, and the forward substitution y_kk = inv(L_kk) b_kk.
// The tile lives in REGISTERS (thread (ty,tx) owns elements (ty+16u, tx+16v)); per pivot only
// column j goes through shared memory (double buffered, one barrier per pivot).  The pivot loop is
// latency bound, so its dependent chain is kept short: the update uses the UNSCALED column and
// 1/d (Newton from an fp32 seed), is applied to all 16 owned elements without predicates (elements
// of finished rows/columns are dead, updating them is harmless), the owner of the next column is
// known at compile time, and the sqrt scaling of the columns happens after the loop.
__device__ void role_potrf(const BaParams &p, const SolverBufs &sb, int kk, double *s_buf) {
  TilePtr Lo = reinterpret_cast<TilePtr>(s_buf);
  TilePtr X = reinterpret_cast<TilePtr>(s_buf + TILE_STRIDE);
  TilePtr Tm = reinterpret_cast<TilePtr>(s_buf + 2 * TILE_STRIDE);
  __shared__ int s_bad;
  __shared__ double s_v[GBA_NB], s_rd[GBA_NB], s_dinv[GBA_NB], s_col[2][2][GBA_NB];
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  double *T = p.sys + (size_t)p.tslot[kk * p.nT + kk] * GBA_NB2;
  if (tid == 0) s_bad = 0;
  double *b = rhs_ptr(p) + kk * GBA_NB;
  const double b_pre = (tid < GBA_NB) ? __ldcg(b + tid) : 0.0;  // rhs block, needed only at the end
  double a[4][4];
#pragma unroll
  for (int u = 0; u < 4; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      const int r = ty + 16 * u, c = tx + 16 * v;
      a[u][v] = (c <= r) ? __ldcg(T + r * GBA_NB + c) : 0.0;
    }
  for (int i = tid; i < GBA_NB * (GBA_NB + 1); i += 256) {
    (&Lo[0][0])[i] = 0.0;
    (&X[0][0])[i] = 0.0;
  }
  // Two pivots per barrier.  The owners publish columns j and j+1 as they stand after pivots < j;
  // every thread derives the pivot-j update of column j+1 itself (B' = B - A * A[j+1] / d0) and
  // applies both rank-1 updates.  1/d1 = d0 / (d0 * B[j+1] - A[j+1]^2) does not wait for 1/d0.
  if (tx < 2) {  // columns 0, 1 -> buffer 0
#pragma unroll
    for (int u = 0; u < 4; u++) s_col[0][tx][ty + 16 * u] = a[u][0];
  }
  __syncthreads();
  bool bad = false;
#pragma unroll
  for (int v0 = 0; v0 < 4; v0++) {  // columns 16 v0 .. 16 v0 + 15
    for (int jj = 0; jj < 16; jj += 2) {
      const int j = 16 * v0 + jj;
      const double *cA = s_col[(j >> 1) & 1][0], *cB = s_col[(j >> 1) & 1][1];
      const double d0 = cA[j], e = cA[j + 1], f = cB[j + 1];
      // rows / columns below 16 v0 are finished: only the sub-blocks u, v >= v0 are still live
      double arA[4], acA[4], arB[4], acB[4];
#pragma unroll
      for (int u = v0; u < 4; u++) {
        arA[u] = cA[ty + 16 * u];
        acA[u] = cA[tx + 16 * u];
        arB[u] = cB[ty + 16 * u];
        acB[u] = cB[tx + 16 * u];
      }
      const double lA = (tid < GBA_NB) ? cA[tid] : 0.0, lB = (tid < GBA_NB) ? cB[tid] : 0.0;
      const double t = fma(d0, f, -e * e);  // d0 * d1
      const double dinv0 = fast_rcp(d0), rt = fast_rcp(t);
      const double m = e * dinv0;           // L[j+1][j] / L[j][j], unscaled form
      const double d1 = fma(-e, m, f);
      const double dinv1 = d0 * rt;
      if (!(d0 > 0.0) || isinf(d0) || !(d1 > 0.0) || isinf(d1)) bad = true;  // SimplicialLLT: NumericalIssue
#pragma unroll
      for (int u = v0; u < 4; u++) {
        arB[u] = fma(-arA[u], m, arB[u]);
        acB[u] = fma(-acA[u], m, acB[u]);
      }
      if (tid < GBA_NB) {  // unscaled columns of L, scaled after the loop
        Lo[tid][j] = (tid >= j) ? lA : 0.0;
        Lo[tid][j + 1] = (tid >= j + 1) ? fma(-lA, m, lB) : 0.0;
      }
      if (tid == 0) {  // the pivots themselves; exact scaling after the loop
        s_dinv[j] = d0;
        s_dinv[j + 1] = d1;
      }
#pragma unroll
      for (int u = v0; u < 4; u++) {
        const double ra = arA[u] * dinv0, rb = arB[u] * dinv1;
#pragma unroll
        for (int v = v0; v < 4; v++) a[u][v] = fma(-rb, acB[v], fma(-ra, acA[v], a[u][v]));
      }
      // owners of columns j+2, j+3 publish them into the other buffer
      if (jj < 14) {
        if (tx == jj + 2 || tx == jj + 3) {
#pragma unroll
          for (int u = 0; u < 4; u++) s_col[((j >> 1) + 1) & 1][tx - jj - 2][ty + 16 * u] = a[u][v0];
        }
      } else if (v0 < 3) {
        if (tx < 2) {
#pragma unroll
          for (int u = 0; u < 4; u++) s_col[((j >> 1) + 1) & 1][tx][ty + 16 * u] = a[u][(v0 + 1) & 3];
        }
      }
      __syncthreads();
    }
  }
  if (bad && tid == 0) s_bad = 1;
  __syncthreads();
  if (s_bad) {
    if (tid == 0) *sb.fail = 1;
    return;
  }
  // scale the columns: L[:,j] = col_j / sqrt(d_j);  s_rd[j] = 1 / L[j][j] = sqrt(dinv_j)
  if (tid < GBA_NB) s_rd[tid] = rsqrt(s_dinv[tid]);
  __syncthreads();
  for (int i = tid; i < GBA_NB2; i += 256) {
    const int r = i >> 6, c = i & 63;
    Lo[r][c] *= s_rd[c];
  }
  __syncthreads();
  // ---- X = inv(L), blocked by 16 -----------------------------------------------------------
  // level 0: the four 16x16 diagonal blocks.  One thread per column, right-looking with the
  // column's 16 partial results in registers: per step one multiply and 15 INDEPENDENT fmas,
  // instead of a dot product whose dependent chain grows with the row.
  if (tid < GBA_NB) {
    const int c = tid, base = (c >> 4) * 16, cc = c - base;
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = (i == cc) ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < 16; j++) {
      const double xj = acc[j] * s_rd[base + j];
      X[base + j][c] = xj;
#pragma unroll
      for (int i = j + 1; i < 16; i++) acc[i] = fma(-Lo[base + i][base + j], xj, acc[i]);
    }
  }
  __syncthreads();
  // block rows I = 1..3: T[I][J] = sum_{K=J}^{I-1} L[I][K] X[K][J];  X[I][J] = -X[I][I] T[I][J].
  // Each thread owns up to 3 outputs (same row-in-block rr, columns cl, cl+16, cl+32) and advances
  // them together with 2 partial sums each, so that 6 independent fma chains are in flight.
#pragma unroll
  for (int I = 1; I < 4; I++) {
    const int rr = tid >> 4, cl = tid & 15, r = 16 * I + rr;
    double s0[3] = {0.0, 0.0, 0.0}, s1[3] = {0.0, 0.0, 0.0}, s2[3] = {0.0, 0.0, 0.0}, s3[3] = {0.0, 0.0, 0.0};
#pragma unroll 2
    for (int m = 0; m < 16 * I; m += 4) {
      const double l0 = Lo[r][m], l1 = Lo[r][m + 1], l2 = Lo[r][m + 2], l3 = Lo[r][m + 3];
#pragma unroll
      for (int q = 0; q < 3; q++)
        if (q < I) {  // X[m][c] is zero above its block's diagonal, so starting at m = 0 is harmless
          s0[q] = fma(l0, X[m][cl + 16 * q], s0[q]);
          s1[q] = fma(l1, X[m + 1][cl + 16 * q], s1[q]);
          s2[q] = fma(l2, X[m + 2][cl + 16 * q], s2[q]);
          s3[q] = fma(l3, X[m + 3][cl + 16 * q], s3[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < 3; q++)
      if (q < I) Tm[rr][cl + 16 * q] = (s0[q] + s1[q]) + (s2[q] + s3[q]);
    __syncthreads();
    double t0[3] = {0.0, 0.0, 0.0}, t1[3] = {0.0, 0.0, 0.0}, t2[3] = {0.0, 0.0, 0.0}, t3[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int m = 0; m < 16; m += 4) {  // X[I][I] is lower triangular: entries right of the diagonal are 0
      const double x0 = X[r][16 * I + m], x1 = X[r][16 * I + m + 1], x2 = X[r][16 * I + m + 2],
                   x3 = X[r][16 * I + m + 3];
#pragma unroll
      for (int q = 0; q < 3; q++)
        if (q < I) {
          t0[q] = fma(x0, Tm[m][cl + 16 * q], t0[q]);
          t1[q] = fma(x1, Tm[m + 1][cl + 16 * q], t1[q]);
          t2[q] = fma(x2, Tm[m + 2][cl + 16 * q], t2[q]);
          t3[q] = fma(x3, Tm[m + 3][cl + 16 * q], t3[q]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 3; q++)
      if (q < I) X[r][cl + 16 * q] = -((t0[q] + t1[q]) + (t2[q] + t3[q]));
    __syncthreads();
  }
  double *XI = sb.invL + (size_t)kk * GBA_NB2;
  for (int i = tid; i < GBA_NB2; i += 256) {
    const int r = i >> 6, c = i & 63;
    T[i] = Lo[r][c];
    XI[i] = X[r][c];
  }
  // forward substitution of the rhs block
  if (tid < GBA_NB) s_v[tid] = b_pre;
  __syncthreads();
  {
    const int row = tid >> 2, part = tid & 3;
    double sa = 0.0, sb_ = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
    for (int m = part * 16; m < part * 16 + 16; m += 4) {
      sa = fma(X[row][m], s_v[m], sa);
      sb_ = fma(X[row][m + 1], s_v[m + 1], sb_);
      sc = fma(X[row][m + 2], s_v[m + 2], sc);
      sd = fma(X[row][m + 3], s_v[m + 3], sd);
    }
    double s = (sa + sb_) + (sc + sd);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (part == 0) b[row] = s;
  }
}

// Rows [q*SLAB, (q+1)*SLAB) of L[i,kk] = A[i,kk] * inv(L_kk)^T, then b_i -= L[i,kk] y_kk there.
// Thread (ty, tx) computes outputs (ty + 16u, tx + 16v): a warp then reads 16 consecutive rows of
// the 65-double-pitch tile per operand, which is bank-conflict free for 8-byte words.
template <int SLAB>
__device__ void role_trsm(const BaParams &p, const SolverBufs &sb, int kk, int a, int q, int atomic_from,
                          double *s_buf) {
  TilePtr X = reinterpret_cast<TilePtr>(s_buf);
  TilePtr A = reinterpret_cast<TilePtr>(s_buf + TILE_STRIDE);
  __shared__ double s_y[GBA_NB];
  constexpr int RPT = SLAB / 16;  // rows per thread
  const int i = sb.colrows[kk * p.nT + a], tid = threadIdx.x, rb = q * SLAB, ty = tid >> 4, tx = tid & 15;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2 + rb * GBA_NB;
  load_rows<GBA_NB>(X, sb.invL + (size_t)kk * GBA_NB2);
  load_rows<SLAB>(A, T);
  if (tid < GBA_NB) s_y[tid] = __ldcg(rhs_ptr(p) + kk * GBA_NB + tid);
  __syncthreads();
  // KS interleaved partial sums per output keep >= 16 independent fp64 fma chains per thread
  constexpr int KS = (RPT >= 4) ? 1 : 4 / RPT;
  double part_[KS][RPT][4] = {};
#pragma unroll 4
  for (int m = 0; m < GBA_NB; m += KS) {
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      double av[RPT], xv[4];
#pragma unroll
      for (int u = 0; u < RPT; u++) av[u] = A[ty + 16 * u][m + ks];
#pragma unroll
      for (int v = 0; v < 4; v++) xv[v] = X[tx + 16 * v][m + ks];
#pragma unroll
      for (int u = 0; u < RPT; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) part_[ks][u][v] = fma(av[u], xv[v], part_[ks][u][v]);
    }
  }
  double acc[RPT][4];
#pragma unroll
  for (int u = 0; u < RPT; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      double t_ = part_[0][u][v];
#pragma unroll
      for (int ks = 1; ks < KS; ks++) t_ += part_[ks][u][v];
      acc[u][v] = t_;
    }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < RPT; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      T[(ty + 16 * u) * GBA_NB + tx + 16 * v] = acc[u][v];
      A[ty + 16 * u][tx + 16 * v] = acc[u][v];
    }
  __syncthreads();
  if (tid < SLAB * 4) {
    const int row = tid >> 2, part = tid & 3;
    double sa = 0.0, sb_ = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
    for (int m = part * 16; m < part * 16 + 16; m += 4) {
      sa = fma(A[row][m], s_y[m], sa);
      sb_ = fma(A[row][m + 1], s_y[m + 1], sb_);
      sc = fma(A[row][m + 2], s_y[m + 2], sc);
      sd = fma(A[row][m + 3], s_y[m + 3], sd);
    }
    double s = (sa + sb_) + (sc + sd);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (part == 0) {
      // closure rows are updated by several band segments at once
      if (i >= atomic_from) atomicAdd(&rhs_ptr(p)[i * GBA_NB + rb + row], -s);
      else rhs_ptr(p)[i * GBA_NB + rb + row] -= s;
    }
  }
}

// Rows [q*SLAB, (q+1)*SLAB) of A[i,j] -= L[i,kk] * L[j,kk]^T
template <int SLAB>
__device__ void role_syrk(const BaParams &p, const SolverBufs &sb, int kk, int a, int b, int q, int atomic_from,
                          double *s_buf) {
  TilePtr B = reinterpret_cast<TilePtr>(s_buf);
  TilePtr A = reinterpret_cast<TilePtr>(s_buf + TILE_STRIDE);
  constexpr int RPT = SLAB / 16;
  const int i = sb.colrows[kk * p.nT + a], j = sb.colrows[kk * p.nT + b], tid = threadIdx.x, rb = q * SLAB;
  const int ty = tid >> 4, tx = tid & 15;
  const double *Li = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2 + rb * GBA_NB;
  const double *Lj = p.sys + (size_t)p.tslot[j * p.nT + kk] * GBA_NB2;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + j] * GBA_NB2 + rb * GBA_NB;
  load_rows<GBA_NB>(B, Lj);
  load_rows<SLAB>(A, Li);
  // the target values are fetched now so that their latency hides behind the product; tiles of
  // the closure block are shared by concurrently factorised band segments -> atomic adds instead
  const bool shared_tile = j >= atomic_from;
  double t[RPT][4];
#pragma unroll
  for (int u = 0; u < RPT; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) t[u][v] = shared_tile ? 0.0 : __ldcg(T + (ty + 16 * u) * GBA_NB + tx + 16 * v);
  __syncthreads();
  constexpr int KS = (RPT >= 4) ? 1 : 4 / RPT;
  double part_[KS][RPT][4] = {};
#pragma unroll 4
  for (int m = 0; m < GBA_NB; m += KS) {
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      double x[RPT], y[4];
#pragma unroll
      for (int u = 0; u < RPT; u++) x[u] = A[ty + 16 * u][m + ks];
#pragma unroll
      for (int v = 0; v < 4; v++) y[v] = B[tx + 16 * v][m + ks];
#pragma unroll
      for (int u = 0; u < RPT; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) part_[ks][u][v] = fma(x[u], y[v], part_[ks][u][v]);
    }
  }
  double acc[RPT][4];
#pragma unroll
  for (int u = 0; u < RPT; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      double t_ = part_[0][u][v];
#pragma unroll
      for (int ks = 1; ks < KS; ks++) t_ += part_[ks][u][v];
      acc[u][v] = t_;
    }
#pragma unroll
  for (int u = 0; u < RPT; u++)
#pragma unroll
    for (int v = 0; v < 4; v++) {
      if (shared_tile) atomicAdd(&T[(ty + 16 * u) * GBA_NB + tx + 16 * v], -acc[u][v]);
      else T[(ty + 16 * u) * GBA_NB + tx + 16 * v] = t[u][v] - acc[u][v];
    }
}

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Fused tile column kk.  Grid = 1 + SPLIT*max_col + SPLIT*max_col(max_col+1)/2 CTAs:
//   CTA 0             potrf of the diagonal tile (skipped when the previous column's look-ahead
//                     already did it), then raises flagA[kk];
//   trsm CTAs         wait for flagA[kk], solve one slab of one tile below the diagonal, count
//                     themselves into cntB[kk];
//   syrk CTAs         wait until cntB[kk] == SPLIT*cnt, update one slab of one trailing tile.
//   look-ahead        the syrk CTAs of tile (kk+1, kk+1) take a ticket; the last one factorises
//                     that tile at once (potrf of column kk+1 overlaps the rest of this update).
// Roles are handed out by an atomic ticket taken when a CTA starts running, not by blockIdx: a
// CTA that waits therefore only ever waits for CTAs that started before it, so the kernel cannot
// deadlock however the grid is dispatched and however much larger than the GPU it is.
// blockIdx.y selects the band segment: tile column kk = kk0 + y * kk_stride (independent columns).
template <int SLAB>
__global__ void __launch_bounds__(256) k_panel(BaParams p, SolverBufs sb, int kk0, int kk_stride, int max_col,
                                               int atomic_from) {
  extern __shared__ double s_buf[];
  constexpr int SPLIT = GBA_NB / SLAB;
  __shared__ int s_flag, s_bid;
  const int tid = threadIdx.x, kk = kk0 + blockIdx.y * kk_stride;
  if (tid == 0) s_bid = (int)atomicAdd(&sb.ticket[kk], 1u);
  __syncthreads();
  const int b = s_bid;
  const int cnt = sb.colcnt[kk];
  if (b == 0) {
    if (ld_acquire(&sb.flagA[kk]) == 0u) {
      if (!__ldcg(sb.fail)) role_potrf(p, sb, kk, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicExch(&sb.flagA[kk], 1u);
    }
    return;
  }
  if (b <= SPLIT * max_col) {
    const int a = (b - 1) / SPLIT, q = (b - 1) % SPLIT;
    if (a >= cnt) return;
    if (tid == 0)
      while (ld_acquire(&sb.flagA[kk]) == 0u) __nanosleep(20);
    __syncthreads();
    if (!__ldcg(sb.fail)) role_trsm<SLAB>(p, sb, kk, a, q, atomic_from, s_buf);
    __threadfence();
    __syncthreads();
    if (tid == 0) atomicAdd(&sb.cntB[kk], 1u);
    return;
  }
  const int w = b - 1 - SPLIT * max_col, pair = w / SPLIT, q = w % SPLIT;
  int a = 0, rem = pair;
  while (rem > a) {
    rem -= a + 1;
    a++;
  }
  if (a >= cnt) return;
  if (tid == 0)
    while (ld_acquire(&sb.cntB[kk]) < (unsigned)(SPLIT * cnt)) __nanosleep(20);
  __syncthreads();
  const bool failed = __ldcg(sb.fail);
  if (!failed) role_syrk<SLAB>(p, sb, kk, a, rem, q, atomic_from, s_buf);
  // look-ahead: tile (kk+1, kk+1) is complete once its SPLIT slabs are updated
  // (not into the closure block: its tiles are still being updated by the other band segments)
  if (pair == 0 && sb.colrows[kk * p.nT] == kk + 1 && kk + 1 < atomic_from) {
    __threadfence();
    __syncthreads();
    if (tid == 0) s_flag = (atomicAdd(&sb.cntD[kk], 1u) == (unsigned)(SPLIT - 1));
    __syncthreads();
    if (s_flag) {
      __threadfence();
      if (!failed) role_potrf(p, sb, kk + 1, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicExch(&sb.flagA[kk + 1], 1u);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_chain: the WHOLE factorisation in ONE launch, as a dependency graph of tile tasks.
//
// With one launch per column (k_panel) a column costs its critical path -- Cholesky of the diagonal
// tile, one triangular solve, one update -- PLUS the tail of its independent updates and the launch
// boundary: ~42 us per closure column and ~78 us per band step on a B200, of which the path itself
// is ~28.  Here the updates of a column run underneath the critical path of the following ones.
//
// Column sequence: the band segments are independent, so their columns are interleaved (step t of
// every segment, then step t+1, ...), followed by the closure columns in order:
//     seq(s) = (s % nseg) * segT + s / nseg   for s < band = nseg * segT,      seq(s) = s   otherwise.
// Tasks, in ticket order (P = potrf, T = trsm slab of 16 rows, S = update; pair 0 is the tile below the
// diagonal, the one the next potrf waits for: it is updated in 4 slabs by 4 CTAs, the other tiles
// whole).  next(k) is k+1 inside a segment and inside the closure block, the first closure column
// after the last band column of the sequence, nothing at the end of the other segments:
//     P(first column of every segment) | block(seq 0) | block(seq 1) | ...
//     block(k) = T(k,*)  S(k,pair 0)  P(next(k))  S(k,rest)
// A persistent CTA takes the next ticket, waits for the task's inputs, runs it, publishes it:
//     P(k)       waits until tile (k,k) has received every update of the columns before k     -> flagA[k]
//     T(k,a,q)   waits for flagA[k] and for tile (i_a,k) to be fully updated        -> upd[(i_a,k)] += TDONE
//     S(k,a,b,q) waits until both operand tiles (i_a,k), (i_b,k) are solved, and for the updates of earlier
//                columns to the target tile (plain read-modify-write: applied in column order) -> upd[(i_a,i_b)]+=
//                Exception: closure tiles updated from band columns -- every segment updates them,
//                in no particular order -- take atomic adds and wait for nobody.
// The number of updates a tile receives from the columns before k is a popcount over the symbolic
// structure bits (columns of other segments never touch a band tile, so "before k" can simply mean
// index < k).  Every wait is for a task with a smaller ticket, i.e. one that is finished or running
// on a resident CTA, so the graph cannot deadlock whatever the grid size; a watchdog turns a broken
// invariant into a failed factorisation (zero update) instead of a hung GPU.
// rhs blocks are updated with atomics: two columns that do not touch each other may both update the
// same row block.
// ---------------------------------------------------------------------------------------------
#define CHAIN_WATCHDOG (1 << 21)  // polls of >= 32 ns: > 60 ms
#define CHAIN_TDONE (1u << 20)      // upd[tile]: low 20 bits count update slabs, the rest solved slabs
__device__ __forceinline__ void chain_wait(const unsigned *ptr, unsigned target, int *fail) {
  int spins = 0;
  while (ld_acquire(ptr) < target) {
    __nanosleep(32);
    if (++spins > CHAIN_WATCHDOG) {
      *fail = 3;
      break;
    }
  }
}

template <int SLAB>
__global__ void __launch_bounds__(256) k_chain(BaParams p, SolverBufs sb, int nseg, int segT, int total) {
  extern __shared__ double s_buf[];
  constexpr int SPLIT = GBA_NB / SLAB;
  __shared__ int s_task[5];  // role, kk, a, b, q
  __shared__ unsigned s_need[2];
  const int tid = threadIdx.x, nT = p.st->n_tiles;
  const int band = nseg > 1 ? nseg * segT : 0, ninit = nseg > 1 ? nseg : 1;
  const int atomic_from = nseg > 1 ? band : (1 << 30);
  auto seq = [&](int s_) { return s_ < band ? (s_ % nseg) * segT + s_ / nseg : s_; };
  auto next_of = [&](int s_, int kk_) {  // column whose potrf is issued inside block(kk_), or -1
    if (s_ < band) {
      if (s_ / nseg + 1 < segT) return kk_ + 1;
      return (s_ == band - 1 && band < nT) ? band : -1;
    }
    return kk_ + 1 < nT ? kk_ + 1 : -1;
  };
  int spos = 0, base = ninit;  // sequence position and first ticket of its block (thread 0 only)
  for (;;) {
    __syncthreads();  // the previous task's shared memory is dead
    if (tid == 0) {
      const int t = (int)atomicAdd(&sb.ticket[0], 1u);
      int role = -1, kk = 0, a = 0, b = 0, q = 0;
      if (t < ninit) {
        role = 0;
        kk = nseg > 1 ? t * segT : 0;
      } else if (t < total) {
        for (;;) {  // tickets only grow: the walk over the sequence is amortised
          const int col = seq(spos), nx = next_of(spos, col);
          const int cnt = sb.colcnt[col], pairs = cnt * (cnt + 1) / 2, nxt = nx >= 0 ? 1 : 0;
          const int n = SPLIT * cnt + (cnt > 0 ? SPLIT + pairs - 1 : 0) + nxt;
          if (t < base + n) {
            int o = t - base;
            kk = col;
            if (o < SPLIT * cnt) {
              role = 1; a = o / SPLIT; q = o % SPLIT;
            } else {
              o -= SPLIT * cnt;
              if (cnt > 0 && o < SPLIT) {
                role = 2; a = 0; b = 0; q = o;
              } else {
                if (cnt > 0) o -= SPLIT;
                if (nxt && o == 0) {
                  role = 0; kk = nx;
                } else {
                  o -= nxt;
                  int pr = 1 + o;  // whole-tile update
                  a = 0;
                  while (pr > a) { pr -= a + 1; a++; }
                  b = pr;
                  role = 3;
                }
              }
            }
            break;
          }
          base += n;
          spos++;
        }
      }
      s_task[0] = role; s_task[1] = kk; s_task[2] = a; s_task[3] = b; s_task[4] = q;
    }
    __syncthreads();
    const int role = s_task[0], kk = s_task[1], a = s_task[2], b = s_task[3], q = s_task[4];
    if (role < 0) return;
    // tiles whose update counts gate this task: (i, j) and the number of earlier columns touching both
    int ti, tj;
    if (role == 0) ti = tj = kk;
    else if (role == 1) { ti = sb.colrows[kk * p.nT + a]; tj = kk; }
    else { ti = sb.colrows[kk * p.nT + a]; tj = sb.colrows[kk * p.nT + b]; }
    // closure tile updated from a band column: atomic, unordered
    const bool unordered = role >= 2 && tj >= atomic_from && kk < band;
    if (tid < 32) {
      unsigned n = 0;
      if (tid < TS_WORDS && !unordered) {
        unsigned m = sb.tbits[ti * TS_WORDS + tid] & sb.tbits[tj * TS_WORDS + tid];
        const int hi = kk - 32 * tid;  // keep bits c < kk of this word
        if (hi <= 0) m = 0;
        else if (hi < 32) m &= (1u << hi) - 1u;
        n = __popc(m);
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
      if (tid == 0) s_need[0] = n * SPLIT;
    }
    __syncthreads();
    const unsigned need = s_need[0];
    const unsigned *updp = sb.upd + p.tslot[ti * p.nT + tj];
    if (tid == 0) {
      if (role == 1) chain_wait(&sb.flagA[kk], 1u, sb.fail);
      if (role >= 2) {  // both operand tiles solved: their counters carry one CHAIN_TDONE per slab
        chain_wait(sb.upd + p.tslot[ti * p.nT + kk], (unsigned)SPLIT * CHAIN_TDONE, sb.fail);
        chain_wait(sb.upd + p.tslot[tj * p.nT + kk], (unsigned)SPLIT * CHAIN_TDONE, sb.fail);
      }
      chain_wait(updp, need, sb.fail);
    }
    __syncthreads();
    const bool failed = __ldcg(sb.fail) != 0;
    if (role == 0) {
      if (!failed) role_potrf(p, sb, kk, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicExch(&sb.flagA[kk], 1u);
    } else if (role == 1) {
      if (!failed) role_trsm<SLAB>(p, sb, kk, a, q, 0, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), CHAIN_TDONE);
    } else if (role == 2) {
      if (!failed) role_syrk<SLAB>(p, sb, kk, a, b, q, kk < band ? atomic_from : (1 << 30), s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), 1u);
    } else {  // a whole tile at once: a quarter of the tasks, the L[j,kk] operand is loaded once instead of 4 times
      if (!failed) role_syrk<GBA_NB>(p, sb, kk, a, b, 0, kk < band ? atomic_from : (1 << 30), s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), (unsigned)SPLIT);
    }
  }
}

// the same three phases as separate launches (large, near-dense structures)
__global__ void __launch_bounds__(256) k_potrf_diag(BaParams p, SolverBufs sb, int kk) {
  extern __shared__ double s_buf[];
  if (!*sb.fail) role_potrf(p, sb, kk, s_buf);
}
__global__ void __launch_bounds__(256) k_trsm(BaParams p, SolverBufs sb, int kk) {
  extern __shared__ double s_buf[];
  if (*sb.fail || (int)blockIdx.x >= sb.colcnt[kk]) return;
  role_trsm<GBA_NB>(p, sb, kk, blockIdx.x, 0, 1 << 30, s_buf);
}
__global__ void __launch_bounds__(256) k_syrk(BaParams p, SolverBufs sb, int kk) {
  extern __shared__ double s_buf[];
  if (*sb.fail) return;
  int a = 0, rem = blockIdx.x;
  while (rem > a) {
    rem -= a + 1;
    a++;
  }
  if (a >= sb.colcnt[kk]) return;
  role_syrk<GBA_NB>(p, sb, kk, a, rem, 0, 1 << 30, s_buf);
}

// ---------------------------------------------------------------------------------------------
// Backward substitution, one launch per tile row i (from the last to the first): CTA 0 computes
// x_i = inv(L_i)^T y_i, the others push y_k -= L[i,k]^T x_i to the active tiles of row i.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_back_step(BaParams p, SolverBufs sb, int i0, int i_stride) {
  __shared__ double s_x[GBA_NB], s_part[4][GBA_NB];
  // thread = (column of the tile, quarter of its rows): lanes run over columns -> coalesced
  const int b = blockIdx.x, tid = threadIdx.x, col = tid & 63, part = tid >> 6;
  const int i = i0 + blockIdx.y * i_stride;  // blockIdx.y = band segment
  double *rhs = rhs_ptr(p);
  if (__ldcg(sb.fail)) return;
  if (b == 0) {
    const double *XI = sb.invL + (size_t)i * GBA_NB2;
    double l[16];
#pragma unroll
    for (int m = 0; m < 16; m++) l[m] = __ldcg(XI + (part * 16 + m) * GBA_NB + col);
    if (tid < GBA_NB) s_x[tid] = __ldcg(rhs + i * GBA_NB + tid);
    __syncthreads();
    double sa = 0.0, sb_ = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
    for (int m = 0; m < 16; m += 4) {
      sa = fma(l[m], s_x[part * 16 + m], sa);
      sb_ = fma(l[m + 1], s_x[part * 16 + m + 1], sb_);
      sc = fma(l[m + 2], s_x[part * 16 + m + 2], sc);
      sd = fma(l[m + 3], s_x[part * 16 + m + 3], sd);
    }
    s_part[part][col] = (sa + sb_) + (sc + sd);
    __syncthreads();
    if (tid < GBA_NB) rhs[i * GBA_NB + tid] = s_part[0][tid] + s_part[1][tid] + s_part[2][tid] + s_part[3][tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) atomicExch(&sb.flagX[i], 1u);
    return;
  }
  if (b - 1 >= sb.rowcnt[i]) return;
  const int k = sb.rowcols[i * p.nT + b - 1];
  const double *L = p.sys + (size_t)p.tslot[i * p.nT + k] * GBA_NB2;
  // prefetch this thread's slice of the tile while waiting for x_i
  double l[16];
#pragma unroll
  for (int m = 0; m < 16; m++) l[m] = __ldcg(L + (part * 16 + m) * GBA_NB + col);
  if (tid == 0)
    while (ld_acquire(&sb.flagX[i]) == 0u) __nanosleep(20);
  __syncthreads();
  if (tid < GBA_NB) s_x[tid] = __ldcg(rhs + i * GBA_NB + tid);
  __syncthreads();
  double sa = 0.0, sb_ = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
  for (int m = 0; m < 16; m += 4) {
    sa = fma(l[m], s_x[part * 16 + m], sa);
    sb_ = fma(l[m + 1], s_x[part * 16 + m + 1], sb_);
    sc = fma(l[m + 2], s_x[part * 16 + m + 2], sc);
    sd = fma(l[m + 3], s_x[part * 16 + m + 3], sd);
  }
  s_part[part][col] = (sa + sb_) + (sc + sd);
  __syncthreads();
  if (tid < GBA_NB) rhs[k * GBA_NB + tid] -= s_part[0][tid] + s_part[1][tid] + s_part[2][tid] + s_part[3][tid];
}

// dx = (float) x in the caller's pose order, or zero when the factorisation failed
// (droid_kernels.cu:1202-1210, chol.py:13-15; the reference hands dx back as fp32, :1204-1205)
__global__ void k_dx_out(BaParams p, SolverBufs sb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int failed = *sb.fail;
  if (i < 6 * p.P) {
    const int pose = i / 6, r = i - 6 * pose;
    p.dx[i] = failed ? 0.f : (float)rhs_ptr(p)[6 * p.perm[pose] + r];
  }
}
__global__ void k_fail_account(BaParams p, SolverBufs sb) {
  if (*sb.fail) {
    p.st->chol_fail += 1;
    p.st->last_fail = *sb.fail;
    *sb.fail = 0;
  }
}
