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
#include "gba_dmma.cuh"

#define TS_MAXT 512  // max tile rows (6 * P <= 32768)
#define TS_WORDS (TS_MAXT / 32)

struct SolverBufs {
  unsigned int *tbits;    // [nT][TS_WORDS] structure bits, row i bit j (j <= i)
  int *colrows, *colcnt;  // active tiles below the diagonal, per tile column
  int *rowcols, *rowcnt;  // active tiles left of the diagonal, per tile row
  double *invL;           // [nT][NB*NB]
  int *fail;              // factorisation hit a non-positive pivot
  unsigned int *flagA, *cntB, *cntD, *flagX, *ticket;  // per-panel in-kernel synchronisation
  unsigned int *cntY, *ticket2;  // backward substitution: updates received per rhs block, its ticket dispenser
  unsigned int *upd;      // [n_slots] slab updates a tile has received inside k_chain
  unsigned long long *trace;  // optional (GBA_CHAIN_TRACE=1): [0] = record count, then 4 words per task of k_chain
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
//  4. Second level (n_seg >= 4): the separators 0, 2, 4, ... touch disjoint pairs of segments, so once the
//     segments are eliminated they are coupled only to the odd separators and to C.  Each becomes a chain of its
//     own (padded to 32 positions) at the head of the closure block; the chains are factorised side by side like
//     the segments, then the dense rest (odd separators + C) as one chain: seg_tiles + seg2_tiles + dense tile
//     columns in sequence.
// Layout of positions: [segment 0 | pad][segment 1 | pad] ... [separator 0 | pad][separator 2 | pad] ...
//                      [odd separators and C, temporal order].
// ---------------------------------------------------------------------------------------------
#define ORD_NC 8
#define ORD_ALIGN 32  // pose positions; 32 * 6 rows = 3 tiles
__global__ void __launch_bounds__(1024) k_order(BaParams p, int pos_cap, int no_level2) {
  __shared__ int s_cnt[ORD_NC], s_reff[ORD_NC];
  __shared__ int s_best, s_warp[32], s_nseg, s_seglen, s_hb, s_SL, s_nA, s_SLA;
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
      // second level: every other separator is independent of the others once the segments are gone
      const int nA = ns >= 4 ? ns / 2 : 0, SLA = (hb + ORD_ALIGN - 1) / ORD_ALIGN * ORD_ALIGN;
      const int chain = (6 * SL + GBA_NB - 1) / GBA_NB + (nA ? 6 * SLA / GBA_NB : 0) +
                        (6 * (nC + nsep - nA * hb) + GBA_NB - 1) / GBA_NB;
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
    // separators 0, 2, 4, ... (between the segments 2m and 2m + 1) touch disjoint pairs of segments: once the
    // segments are eliminated they are coupled only to their odd neighbours and to C -- they are factorised side
    // by side as a second set of chains, in front of the dense rest (odd separators + C)
    s_SLA = (hb + ORD_ALIGN - 1) / ORD_ALIGN * ORD_ALIGN;
    s_nA = best_seg >= 4 ? best_seg / 2 : 0;
    if (s_nA && best_seg * s_SL + s_nA * s_SLA + nC + (best_seg - 1 - s_nA) * hb > pos_cap) s_nA = 0;
    if (no_level2) s_nA = 0;
  }
  __syncthreads();
  const int bc = s_best, nseg = s_nseg, seglen = s_seglen, hb = s_hb, SL = s_SL, nA = s_nA, SLA = s_SLA;
  const int band_end = nseg * SL;  // first position of the closure block (nseg == 1: == nB)
  const int core0 = band_end + nA * SLA;  // first position of the dense rest of the closure block
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
    int seg = 0, off = 0, clo = inC, posA = -1;
    if (i < P && !inC) {
      const int bidx = bincl - 1;
      if (nseg > 1) {
        seg = bidx / (seglen + hb);
        off = bidx - seg * (seglen + hb);
        if (seg >= nseg) {  // tail beyond the last separator belongs to the last segment
          seg = nseg - 1;
          off = bidx - seg * (seglen + hb);
        }
        if (off >= seglen && seg < nseg - 1) {              // separator number seg
          if (nA && (seg & 1) == 0) posA = band_end + (seg >> 1) * SLA + (off - seglen);
          else clo = 1;
        } else if (off >= SL) clo = 1;                      // overflow of the last segment
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
      const int pos = posA >= 0 ? posA : (clo ? core0 + cincl - 1 : seg * SL + off);
      p.perm[i] = pos;
      p.iperm[pos] = i;
    }
    carry_band = total_band;
    carry_clo = total_clo;
  }
  if (tid == 0) {
    const int npos = core0 + carry_clo;
    p.st->n_closure = carry_clo + nA * hb;
    p.st->n_seg2 = nA;
    p.st->seg2_tiles = nA ? 6 * SLA / GBA_NB : 0;
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
  // slots: row-major over the lower triangle, but the closure x closure tiles (both tile indices >= bandT) are
  // numbered after all the others: together with the rhs that follows the tiles they form ONE contiguous
  // range -- what the ranks of an edge-sharded run exchange (every other tile is produced by one rank alone).
  const int bandT = p.st->n_seg > 1 ? p.st->n_seg * p.st->seg_tiles : 0;
  auto popc_below = [&](int i, int j) {  // structure bits of row i in columns [0, j)
    int c = 0;
    for (int w = 0; w < (j >> 5); w++) c += __popc(bits[i * TS_WORDS + w]);
    if (j & 31) c += __popc(bits[i * TS_WORDS + (j >> 5)] & ((1u << (j & 31)) - 1u));
    return c;
  };
  __shared__ int s_start_cc[TS_MAXT + 1];
  for (int i = tid; i < nT; i += blockDim.x) {
    const int all = popc_below(i, i + 1), nc = popc_below(i, min(bandT, i + 1));
    s_start[i + 1] = nc;            // tiles of row i left of the closure block
    s_start_cc[i + 1] = all - nc;   // its closure x closure tiles
  }
  __syncthreads();
  if (tid == 0) {
    s_start[0] = 0;
    s_start_cc[0] = 0;
    for (int i = 0; i < nT; i++) {
      s_start[i + 1] += s_start[i];
      s_start_cc[i + 1] += s_start_cc[i];
    }
    p.st->cc_slot0 = s_start[nT];
    p.st->n_slots = s_start[nT] + s_start_cc[nT];
  }
  __syncthreads();
  for (int idx = tid; idx < nT * nT; idx += blockDim.x) {
    const int i = idx / nT, j = idx % nT;
    int slot = -1;
    if (j <= i && ((bits[i * TS_WORDS + (j >> 5)] >> (j & 31)) & 1)) {
      if (j < bandT) slot = s_start[i] + popc_below(i, j);
      else slot = s_start[nT] + s_start_cc[i] + popc_below(i, j) - popc_below(i, bandT);
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
// Tile-column roles (256 threads; dynamic shared memory DM_SMEM_BYTES).  The 64 x 64 fp64 tile
// operations run on the tensor pipe (mma.sync.m8n8k4.f64, gba_dmma.cuh); the trsm / update roles
// work on SLAB-row slabs of a tile (SPLIT = 64 / SLAB CTAs per tile) to spread a tile column over
// more SMs while the structure is sparse.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double *rhs_ptr(const BaParams &p) { return p.sys + (size_t)p.st->n_slots * GBA_NB2; }

// Cholesky of the diagonal tile: inverse of its factor into invL[kk] (the factor itself is never needed
// again: the solves and substitutions use the inverse), forward substitution y_kk = inv(L_kk) b_kk.
// A pivot that is not positive and finite raises the failure flag (SimplicialLLT: NumericalIssue).
__device__ void role_potrf(const BaParams &p, const SolverBufs &sb, int kk, bool fused, double *s_buf) {
  DmPotrfSmem &s = *reinterpret_cast<DmPotrfSmem *>(s_buf);
  const double *T = p.sys + (size_t)p.tslot[kk * p.nT + kk] * GBA_NB2;
  // fused: the previous column of the chain has not updated this tile; its tile L[kk,kk-1] is solved (k_chain)
  const double *Lprev = fused ? p.sys + (size_t)p.tslot[kk * p.nT + kk - 1] * GBA_NB2 : nullptr;
  const bool ok = dm_potrf64(T, sb.invL + (size_t)kk * GBA_NB2, rhs_ptr(p) + kk * GBA_NB, s, Lprev);
  if (!ok && threadIdx.x == 0) *sb.fail = 1;
}

// Rows [q*SLAB, (q+1)*SLAB) of L[i,kk] = A[i,kk] * inv(L_kk)^T, then b_i -= L[i,kk] y_kk there.
template <int SLAB>
__device__ void role_trsm(const BaParams &p, const SolverBufs &sb, int kk, int a, int q, int atomic_from,
                          double *s_buf) {
  DmSlabSmem &s = *reinterpret_cast<DmSlabSmem *>(s_buf);
  constexpr int RB = SLAB / 8;
  const int i = sb.colrows[kk * p.nT + a], tid = threadIdx.x, lane = tid & 31, w = tid >> 5, rb0 = q * SLAB;
  const int g = lane >> 2, t = lane & 3;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2 + rb0 * GBA_NB;
  dm_load_rows<GBA_NB>(s.B, sb.invL + (size_t)kk * GBA_NB2);
  dm_load_rows<SLAB>(s.A, T);
  if (tid < GBA_NB) s.y[tid] = __ldcg(rhs_ptr(p) + kk * GBA_NB + tid);
  __syncthreads();
  double acc[RB][2];
#pragma unroll
  for (int rb = 0; rb < RB; rb++) acc[rb][0] = acc[rb][1] = 0.0;
  dm_slab_nt<RB>(acc, s, w, 8 * (w + 1), lane);  // inv(L) is lower triangular: row c ends at column c
  const double y0 = s.y[8 * w + 2 * t], y1 = s.y[8 * w + 2 * t + 1];
#pragma unroll
  for (int rb = 0; rb < RB; rb++) {
    *reinterpret_cast<double2 *>(T + (8 * rb + g) * GBA_NB + 8 * w + 2 * t) = make_double2(acc[rb][0], acc[rb][1]);
    double v = fma(acc[rb][0], y0, acc[rb][1] * y1);  // this warp's 8 columns of (L y)[row]
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    if (t == 0) s.part[w][8 * rb + g] = v;
  }
  __syncthreads();
  if (tid < SLAB) {
    double v = 0.0;
#pragma unroll
    for (int ww = 0; ww < 8; ww++) v += s.part[ww][tid];
    // closure rows are updated by several band segments at once
    if (i >= atomic_from) atomicAdd(&rhs_ptr(p)[i * GBA_NB + rb0 + tid], -v);
    else rhs_ptr(p)[i * GBA_NB + rb0 + tid] -= v;
  }
}

// Rows [q*SLAB, (q+1)*SLAB) of A[i,j] -= L[i,kk] * L[j,kk]^T
template <int SLAB>
__device__ void role_syrk(const BaParams &p, const SolverBufs &sb, int kk, int a, int b, int q, int atomic_from,
                          double *s_buf) {
  DmSlabSmem &s = *reinterpret_cast<DmSlabSmem *>(s_buf);
  constexpr int RB = SLAB / 8;
  const int i = sb.colrows[kk * p.nT + a], j = sb.colrows[kk * p.nT + b], tid = threadIdx.x, rb0 = q * SLAB;
  const int lane = tid & 31, w = tid >> 5, g = lane >> 2, t = lane & 3;
  const double *Li = p.sys + (size_t)p.tslot[i * p.nT + kk] * GBA_NB2 + rb0 * GBA_NB;
  const double *Lj = p.sys + (size_t)p.tslot[j * p.nT + kk] * GBA_NB2;
  double *T = p.sys + (size_t)p.tslot[i * p.nT + j] * GBA_NB2 + rb0 * GBA_NB;
  dm_load_rows<GBA_NB>(s.B, Lj);
  dm_load_rows<SLAB>(s.A, Li);
  // the target values are fetched now so that their latency hides behind the product; tiles of
  // the closure block are shared by concurrently factorised band segments -> atomic adds instead
  const bool shared_tile = j >= atomic_from;
  double2 tv[RB];
#pragma unroll
  for (int rb = 0; rb < RB; rb++)
    tv[rb] = shared_tile ? make_double2(0.0, 0.0)
                         : __ldcg(reinterpret_cast<const double2 *>(T + (8 * rb + g) * GBA_NB + 8 * w + 2 * t));
  __syncthreads();
  double acc[RB][2];
#pragma unroll
  for (int rb = 0; rb < RB; rb++) acc[rb][0] = acc[rb][1] = 0.0;
  dm_slab_nt<RB>(acc, s, w, GBA_NB, lane);
#pragma unroll
  for (int rb = 0; rb < RB; rb++) {
    double *dst = T + (8 * rb + g) * GBA_NB + 8 * w + 2 * t;
    if (shared_tile) {
      atomicAdd(dst, -acc[rb][0]);
      atomicAdd(dst + 1, -acc[rb][1]);
    } else {
      *reinterpret_cast<double2 *>(dst) = make_double2(tv[rb].x - acc[rb][0], tv[rb].y - acc[rb][1]);
    }
  }
}

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
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
// every segment, then step t+1, ...), followed by the second-level separators interleaved the same way, followed
// by the dense rest of the closure block in order:
//     seq(s) = (s % nseg) * segT + s / nseg   for s < band = nseg * segT,   likewise from band on for the nseg2
//     separators of segT2 columns,   seq(s) = s   otherwise.
// Tasks, in ticket order (P = potrf, T = trsm slab of 16 rows, S = update; pair 0 is the tile below the
// diagonal, the one the next potrf waits for: it is updated in 4 slabs by 4 CTAs, the other tiles
// whole).  next(k) is k+1 inside a segment and inside the closure block, the first closure column
// after the last band column of the sequence, nothing at the end of the other segments:
//     P(first column of every segment) | block(seq 0) | block(seq 1) | ...
//     block(k) = T(k,*)  S(k,pair 0)  P(next(k))  S(k,rest)
// split over two ticket queues (k_chain): SPINE = T(k,0,*) S(k,pair 0) P(next(k)), BULK = the rest.  When the next
// column of the same chain sits right below the diagonal, S(k,pair 0) does not exist: P(k+1) waits for T(k,0,*) and
// subtracts L[k+1,k] L[k+1,k]^T from its tile itself (FUSED, dm_potrf64).
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
__device__ __forceinline__ unsigned long long chain_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define CHAIN_TRACE_CAP 65536
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
__global__ void __launch_bounds__(256, 2) k_chain(BaParams p, SolverBufs sb, int nseg, int segT, int total_spine,
                                               int total_bulk, int n_spine, int phase, int seg_lo, int seg_n,
                                               int nseg2, int segT2) {
  extern __shared__ double s_buf[];
  constexpr int SPLIT = GBA_NB / SLAB;
  __shared__ int s_task[5];  // role, kk, a, b, q
  __shared__ unsigned s_need[2];
  const int tid = threadIdx.x, nT = p.st->n_tiles;
  // phase 0: the whole factorisation.  An edge-sharded run splits it: phase 1 = the columns of this rank's band
  // segments [seg_lo, seg_lo + seg_n) only (their updates of closure tiles are unordered atomics), phase 2 = the
  // closure block alone, after the ranks have summed it (updates of band columns are then already in the tiles:
  // only the columns >= band count as predecessors).
  // Three stages of columns: the band segments [0, band) side by side, then -- second dissection level -- nseg2
  // separators [band, band2) side by side (segT2 tile columns each, k_order), then the dense rest [band2, nT) as
  // one chain.  The closure block of the sharded solve is everything from `band` on.
  const int band = nseg > 1 ? nseg * segT : 0;
  const int band2 = band + nseg2 * segT2;
  const int nsl = phase == 1 ? seg_n : (phase == 2 ? 0 : (nseg > 1 ? nseg : 0));  // band segments in play
  const int seg0 = phase == 1 ? seg_lo : 0;
  const int nsA = phase == 1 ? 0 : nseg2;                                           // second-level chains in play
  const int nband = nsl * segT, nbandA = nsA * segT2;                              // sequence positions of the two
  const int clo = phase == 1 ? 0 : nT - band2;                                      // ... and of the dense rest
  const int npos = nband + nbandA + clo;
  const int ninit = nsl > 0 ? nsl : (nsA > 0 ? nsA : 1);
  const int minbit = phase == 2 ? band : 0;
  // a tile beyond the end of the stage that updates it is shared by that stage's chains: atomic adds, no order
  auto atomic_from_of = [&](int col) { return col < band ? band : (col < band2 ? band2 : (1 << 30)); };
  // Two ticket queues.  SPINE: the tasks every following column waits for -- P(k), the solve of the tile below
  // the diagonal and the update of the next diagonal tile -- served by the first n_spine CTAs only, so that
  // they never queue behind a wave of bulk updates.  BULK: everything else.  Inside each queue tickets are
  // handed out in sequence order; a waiting task only waits for tasks that precede it in the combined order,
  // and the earliest unfinished task of the combined order is always the head of its own queue, hence taken
  // and runnable: no deadlock for any grid size >= n_spine + 1.
  // n_spine < 0: the grid fills the machine with exactly two CTAs per SM, and the spine CTAs are the PAIRS living on
  // the SMs 0 .. -n_spine/2 - 1: a factorisation then never shares its SM (issue slots, DMMA pipe) with a bulk
  // update running flat out, only with another spine task (12.4 us alone against 15-18 us beside a bulk CTA).
  bool spine;
  if (n_spine < 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    spine = (int)smid < (-n_spine) / 2;
  } else {
    spine = (int)blockIdx.x < n_spine;
  }
  unsigned *ticket = spine ? &sb.ticket[0] : &sb.ticket[1];
  (void)total_spine; (void)total_bulk;  // upper bounds, used by the host to size the grid: the lists end with the sequence
  auto seq = [&](int s_) {
    if (s_ < nband) return (seg0 + s_ % nsl) * segT + s_ / nsl;
    s_ -= nband;
    if (s_ < nbandA) return band + (s_ % nsA) * segT2 + s_ / nsA;
    return band2 + (s_ - nbandA);
  };
  // the potrf(s) issued inside block(kk_): the next column of the same chain (n = 1, chain = true), or -- at the
  // last sequence position of a stage -- the first columns of all chains of the next stage (first, stride, n)
  struct Next { int first, stride, n; bool chain; };
  auto stage_heads = [&](int from_stage) {  // heads of the first non-empty stage after from_stage (0 band, 1 level 2)
    Next r = {-1, 0, 0, false};
    if (from_stage == 0 && nbandA > 0) { r.first = band; r.stride = segT2; r.n = nsA; }
    else if (clo > 0) { r.first = band2; r.stride = 0; r.n = 1; }
    return r;
  };
  auto next_of = [&](int s_, int kk_) {
    Next r = {-1, 0, 0, false};
    if (s_ < nband) {
      if (s_ / nsl + 1 < segT) { r.first = kk_ + 1; r.n = 1; r.chain = true; }
      else if (s_ == nband - 1) r = stage_heads(0);
    } else if (s_ < nband + nbandA) {
      const int sa = s_ - nband;
      if (sa / nsA + 1 < segT2) { r.first = kk_ + 1; r.n = 1; r.chain = true; }
      else if (sa == nbandA - 1) r = stage_heads(1);
    } else if (kk_ + 1 < nT) { r.first = kk_ + 1; r.n = 1; r.chain = true; }
    return r;
  };
  int spos = 0, base = spine ? ninit : 0;  // sequence position and first ticket of its block (thread 0 only)
  for (;;) {
    __syncthreads();  // the previous task's shared memory is dead
    if (tid == 0) {
      const int t = (int)atomicAdd(ticket, 1u);
      int role = -1, kk = 0, a = 0, b = 0, q = 0;
      if (spine && t < ninit) {
        role = 0;
        kk = nsl > 0 ? (seg0 + t) * segT : (nsA > 0 ? band + t * segT2 : band2);
      } else {
        while (spos < npos) {  // tickets only grow: the walk over the sequence is amortised
          const int col = seq(spos);
          const Next nx = next_of(spos, col);
          const int cnt = sb.colcnt[col], pairs = cnt * (cnt + 1) / 2;
          // FUSED: the next column of the same chain sits right below the diagonal -- its factorisation applies
          // this column's update of its diagonal tile itself (dm_potrf64, Lprev) instead of waiting for an update task
          const bool fz = cnt > 0 && nx.chain && sb.colrows[col * p.nT] == nx.first;
          // spine block(k): T(k, 0, *)  [S(k, pair 0, *) unless fused]  P(next(k)) ...;
          // bulk block(k): T(k, a >= 1, *)  S(k, other pairs)
          const int ncrit = cnt > 0 ? (fz ? SPLIT : 2 * SPLIT) : 0;
          const int n = spine ? ncrit + nx.n : (cnt > 0 ? SPLIT * (cnt - 1) + pairs - 1 : 0);
          if (t < base + n) {
            int o = t - base;
            kk = col;
            if (spine) {
              if (cnt > 0 && o < SPLIT) {
                role = 1; a = 0; q = o;
              } else if (cnt > 0 && !fz && o < 2 * SPLIT) {
                role = 2; a = 0; b = 0; q = o - SPLIT;
              } else {
                role = 0; kk = nx.first + (o - ncrit) * nx.stride; b = fz ? 1 : 0;
              }
            } else {
              if (o < SPLIT * (cnt - 1)) {
                role = 1; a = 1 + o / SPLIT; q = o % SPLIT;
              } else {
                int pr = 1 + (o - SPLIT * (cnt - 1));  // whole-tile update, pairs 1 .. pairs-1
                a = 0;
                while (pr > a) { pr -= a + 1; a++; }
                b = pr;
                role = 3;
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
    unsigned long long tr_take = 0, tr_run = 0;
    if (sb.trace && tid == 0) tr_take = chain_now();
    // tiles whose update counts gate this task: (i, j) and the number of earlier columns touching both
    int ti, tj;
    if (role == 0) ti = tj = kk;
    else if (role == 1) { ti = sb.colrows[kk * p.nT + a]; tj = kk; }
    else { ti = sb.colrows[kk * p.nT + a]; tj = sb.colrows[kk * p.nT + b]; }
    // closure tile updated from a band column: atomic, unordered
    const int atomic_from = atomic_from_of(kk);
    const bool unordered = role >= 2 && tj >= atomic_from;
    if (tid < 32) {
      unsigned n = 0;
      if (tid < TS_WORDS && !unordered) {
        unsigned m = sb.tbits[ti * TS_WORDS + tid] & sb.tbits[tj * TS_WORDS + tid];
        const int hi = kk - 32 * tid, lo = minbit - 32 * tid;  // keep bits minbit <= c < kk of this word
        if (hi <= 0) m = 0;
        else if (hi < 32) m &= (1u << hi) - 1u;
        if (lo >= 32) m = 0;
        else if (lo > 0) m &= ~((1u << lo) - 1u);
        n = __popc(m);
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
      if (tid == 0) s_need[0] = (n - ((role == 0 && b) ? 1u : 0u)) * SPLIT;  // fused: column kk - 1 sends no update
    }
    __syncthreads();
    const unsigned need = s_need[0];
    const unsigned *updp = sb.upd + p.tslot[ti * p.nT + tj];
    if (tid == 0) {
      if (role == 1) chain_wait(&sb.flagA[kk], 1u, sb.fail);
      if (role == 0 && b) chain_wait(sb.upd + p.tslot[kk * p.nT + kk - 1], (unsigned)SPLIT * CHAIN_TDONE, sb.fail);
      if (role >= 2) {  // both operand tiles solved: their counters carry one CHAIN_TDONE per slab
        chain_wait(sb.upd + p.tslot[ti * p.nT + kk], (unsigned)SPLIT * CHAIN_TDONE, sb.fail);
        chain_wait(sb.upd + p.tslot[tj * p.nT + kk], (unsigned)SPLIT * CHAIN_TDONE, sb.fail);
      }
      chain_wait(updp, need, sb.fail);
      if (sb.trace) tr_run = chain_now();
    }
    __syncthreads();
    const bool failed = __ldcg(sb.fail) != 0;
    if (role == 0) {
      if (!failed) role_potrf(p, sb, kk, b != 0, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicExch(&sb.flagA[kk], 1u);
    } else if (role == 1) {
      if (!failed) role_trsm<SLAB>(p, sb, kk, a, q, 0, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), CHAIN_TDONE);
    } else if (role == 2) {
      if (!failed) role_syrk<SLAB>(p, sb, kk, a, b, q, atomic_from, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), 1u);
    } else {  // a whole tile at once: a quarter of the tasks, the L[j,kk] operand is loaded once instead of 4 times
      if (!failed) role_syrk<GBA_NB>(p, sb, kk, a, b, 0, atomic_from, s_buf);
      __threadfence();
      __syncthreads();
      if (tid == 0) atomicAdd(const_cast<unsigned *>(updp), (unsigned)SPLIT);
    }
    if (sb.trace && tid == 0) {
      const unsigned long long n = atomicAdd(sb.trace, 1ull);
      if (n < CHAIN_TRACE_CAP) {
        unsigned long long *r = sb.trace + 1 + 4 * n;
        unsigned smid;
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        r[0] = ((unsigned long long)role << 56) | ((unsigned long long)kk << 40) | ((unsigned long long)a << 28) | ((unsigned long long)b << 16) | ((unsigned long long)q << 12) | (smid & 0xfff);
        r[1] = tr_take; r[2] = tr_run; r[3] = chain_now();
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward substitution L^T x = y in ONE launch: a task per active tile (i, k), k < i, handed out by
// ticket in order of descending tile row i.  Task (i, k) waits for x_i (flagX[i]), subtracts
// L[i,k]^T x_i from y_k (fp64 atomics: rows that do not touch each other update the same block
// concurrently) and counts itself into cntY[k]; whoever delivers the LAST update of y_k computes
// x_k = inv(L_k)^T y_k on the spot and raises flagX[k] -- one hop per tile row on the critical path.
// Rows that receive no update (no active tile below their diagonal tile: the last row) are seeded
// by the first tickets.  Every wait is for a task with a smaller ticket (rows > i come first), so
// the kernel cannot deadlock whatever the grid size; the watchdog of k_chain applies.
// Replaces solver.solve(b)'s backward half (droid_kernels.cu:1203) and 37 launches per iteration.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void back_solve_x(const BaParams &p, const SolverBufs &sb, int i, const double (&xi)[16],
                                             double *s_x, double (*s_part)[GBA_NB]) {
  // x_i = inv(L_i)^T y_i; thread = (column col of inv(L), quarter `part` of its rows), xi = its slice of inv(L_i)
  const int tid = threadIdx.x, col = tid & 63, part = tid >> 6;
  double *rhs = rhs_ptr(p);
  if (tid < GBA_NB) s_x[tid] = __ldcg(rhs + i * GBA_NB + tid);
  __syncthreads();
  double sa = 0.0, sb_ = 0.0, sc = 0.0, sd = 0.0;
#pragma unroll
  for (int m = 0; m < 16; m += 4) {
    sa = fma(xi[m], s_x[part * 16 + m], sa);
    sb_ = fma(xi[m + 1], s_x[part * 16 + m + 1], sb_);
    sc = fma(xi[m + 2], s_x[part * 16 + m + 2], sc);
    sd = fma(xi[m + 3], s_x[part * 16 + m + 3], sd);
  }
  s_part[part][col] = (sa + sb_) + (sc + sd);
  __syncthreads();
  if (tid < GBA_NB) rhs[i * GBA_NB + tid] = s_part[0][tid] + s_part[1][tid] + s_part[2][tid] + s_part[3][tid];
  __threadfence();
  __syncthreads();
  if (tid == 0) atomicExch(&sb.flagX[i], 1u);
}

// seg_n >= 1 (edge-sharded run): only the closure rows and the rows of this rank's band segments are solved here.
__global__ void __launch_bounds__(256) k_back(BaParams p, SolverBufs sb, int total, int seg_lo, int seg_n, int nseg,
                                              int segT) {
  __shared__ double s_x[GBA_NB], s_part[4][GBA_NB];
  __shared__ int s_task[3];  // role (0 seed, 1 tile, -1 done), i, k
  const int tid = threadIdx.x, col = tid & 63, part = tid >> 6, nT = p.st->n_tiles;
  if (__ldcg(sb.fail)) return;
  const int band = nseg * segT;
  auto own = [&](int tile) { return seg_n < 0 || tile >= band || (tile / segT >= seg_lo && tile / segT < seg_lo + seg_n); };
  int row = nT - 1, base = 0, nseed = -1;  // thread 0: walk over the rows, amortised (tickets only grow)
  for (;;) {
    __syncthreads();
    if (tid == 0) {
      const int tk = (int)atomicAdd(sb.ticket2, 1u);
      int role = -1, i = 0, k = 0;
      if (nseed < 0) {
        nseed = 0;
        for (int r = 0; r < nT; r++) nseed += (sb.colcnt[r] == 0);
        base = nseed;
      }
      if (tk < nseed) {  // the tk-th row (from the end) without updates
        int seen = 0;
        for (int r = nT - 1; r >= 0; r--)
          if (sb.colcnt[r] == 0 && seen++ == tk) {
            role = 0;
            i = r;
            break;
          }
      } else if (tk < total) {
        while (row >= 0 && tk >= base + sb.rowcnt[row]) base += sb.rowcnt[row--];
        if (row >= 0) {
          role = 1;
          i = row;
          k = sb.rowcols[row * p.nT + (tk - base)];
        }
      }
      s_task[0] = role; s_task[1] = i; s_task[2] = k;
    }
    __syncthreads();
    const int role = s_task[0], i = s_task[1], k = s_task[2];
    if (role < 0) return;
    if (!own(role == 0 ? i : k)) continue;  // another rank's band segment
    double l[16];
    if (role == 0) {
      const double *XI = sb.invL + (size_t)i * GBA_NB2;
#pragma unroll
      for (int m = 0; m < 16; m++) l[m] = __ldcg(XI + (part * 16 + m) * GBA_NB + col);
      back_solve_x(p, sb, i, l, s_x, s_part);
      continue;
    }
    // this thread's slices of L[i,k] and of inv(L_k), requested before the wait
    const double *L = p.sys + (size_t)p.tslot[i * p.nT + k] * GBA_NB2, *XK = sb.invL + (size_t)k * GBA_NB2;
    double xk[16];
#pragma unroll
    for (int m = 0; m < 16; m++) l[m] = __ldcg(L + (part * 16 + m) * GBA_NB + col);
#pragma unroll
    for (int m = 0; m < 16; m++) xk[m] = __ldcg(XK + (part * 16 + m) * GBA_NB + col);
    if (tid == 0) chain_wait(&sb.flagX[i], 1u, sb.fail);
    __syncthreads();
    if (__ldcg(sb.fail)) return;
    if (tid < GBA_NB) s_x[tid] = __ldcg(rhs_ptr(p) + i * GBA_NB + tid);
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
    if (tid < GBA_NB)
      atomicAdd(rhs_ptr(p) + k * GBA_NB + tid, -(s_part[0][tid] + s_part[1][tid] + s_part[2][tid] + s_part[3][tid]));
    __threadfence();
    __syncthreads();
    if (tid == 0) s_task[0] = (atomicAdd(&sb.cntY[k], 1u) == (unsigned)(sb.colcnt[k] - 1));
    __syncthreads();
    if (s_task[0]) {  // y_k is complete
      __threadfence();
      back_solve_x(p, sb, k, xk, s_x, s_part);
    }
  }
}

// dx = (float) x in the caller's pose order, or zero when the factorisation failed
// (droid_kernels.cu:1202-1210, chol.py:13-15; the reference hands dx back as fp32, :1204-1205)
// (edge-sharded run: poses of another rank's band segments get a zero update here; their owner moves them)
__global__ void k_dx_out(BaParams p, SolverBufs sb, int seg_lo, int seg_n, int nseg, int segT) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int failed = *sb.fail;
  if (i < 6 * p.P) {
    const int pose = i / 6, r = i - 6 * pose, pos = p.perm[pose], tile = 6 * pos / GBA_NB;
    const bool own = seg_n < 0 || tile >= nseg * segT || (tile / segT >= seg_lo && tile / segT < seg_lo + seg_n);
    p.dx[i] = (failed || !own) ? 0.f : (float)rhs_ptr(p)[6 * pos + r];
  }
}
__global__ void k_fail_account(BaParams p, SolverBufs sb) {
  if (*sb.fail) {
    p.st->chol_fail += 1;
    p.st->last_fail = *sb.fail;
    *sb.fail = 0;
  }
}
