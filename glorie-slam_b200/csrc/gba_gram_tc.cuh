// gba_gram_tc.cuh -- Schur Gram G = A A^T on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// Same contract as k_gram (gba_kernels.cuh): for depth keyframe k with out-edges n_1..n_d, the rows
// a_n[c][pix] = sqrt(Q) * sum_{u,v} w J_z J_j[c] (c = 0..5) plus the row sqrt(Q) * w_k are rebuilt
// tile by tile in shared memory and G = A A^T lands in p.gram + goff[k] (pitch Rp).  Replaces the
// reference's Eii/Eij materialisation + EEt6x6/Ev6x1 (droid_kernels.cu:340-341,373-374,1001-1093,
// 1222-1296), like k_gram.
//
// What is different from k_gram: the 6 deg + 1 rows of a keyframe (<= 128, i.e. deg <= 21) are one
// M = 128 operand tile and the products are tcgen05.mma.kind::tf32 instructions issued by a single
// thread into fp32 accumulators in tensor memory.  All the CUDA cores do is rebuild the rows.
//
// Accuracy.  (1) 3xTF32: a = hi + lo with hi = rna_tf32(a); G = hi hi^T + (hi lo^T + lo hi^T), the
// lo lo^T term is 2^-22 relative.  (2) The tensor core truncates (round toward zero) every time it
// adds a K = 8 block of products to the accumulator -- measured here as a systematic shrink of about
// 3e-8 x (number of accumulations); one accumulator over 3072 pixels would be 1e-5 low, which the
// Schur complement A - S amplifies.  So the big term hi hi^T and the small cross terms go to
// different accumulators (D1, D2: truncating D2 is harmless), and every GT_GROUP tiles (16
// accumulations) the accumulators are drained into fp32 registers with ordinary round-to-nearest
// adds while the MMAs continue into a second accumulator pair (ping-pong): 2 x (D1 + D2) x 128
// columns = all 512 TMEM columns, hence one CTA per SM.
//
// Shared-memory operand layout (both hi and lo): K-major, no swizzle, 8 x 16 B core matrices
//   byte(r, k) = (r % 8) * 16 + (r / 8) * SBO + (k % 4) * 4 + (k / 4) * LBO,   SBO = 128,
//   LBO = 16 * 128 + 16 = 2064: the 16 B pad makes the 32 pixel lanes of a producer warp, which all
//   store the same row r at k = lane, fall into 32 different banks.
// Pipeline (16 warps, one CTA per SM):
//   warp 14, loader: 16-byte cp.async of the raw inputs -- per edge the two weight rows (the E rows carry no
//     residual, so the targets are not read here), per
//     keyframe disparity, Q, w -- 64 pixels (one "slot") at a time into a GT_IS-deep ring; each lane's
//     cp.async.mbarrier.arrive fires on in_full[] when its copies have landed.  ~45 KB in flight per
//     SM without a single register.
//   warps 0-13, producers: per slot a warp rebuilds the rows of one or two edges (which warps get two
//     rotates with the slot) for 2 x 32 pixels, stores hi/lo into the operand ring, arrives on
//     full[stage] of both tiles and releases the input slot on in_empty[].  No CTA-wide barrier.
//   warp 15, MMA issue: waits on full, issues 4 k-steps x 3 MMAs, tcgen05.commit -> empty[stage]; at a
//     group end commits -> accfull[set].
//   Producers poll accfull after every slot and inside every wait loop (nobody can owe a drain while
//   blocked), drain their 32 x 32 chunk of the set into registers and arrive on accfree[set].  Warp w
//   owns TMEM lanes 32 (w % 4) .. + 31 and the column chunk (w % 4) + (w / 4): the 10 chunks of the
//   block-upper triangle all fall on warps 0-12.
// Three rules keep the mbarrier phase-parity waits sound: every producer warp takes part in both rings
// for every slot even when it has no edge there (nobody can be lapped), every probe is made
// warp-uniform by a vote before anything warp-collective (tcgen05.ld) depends on it, and every wait
// carries a watchdog (GBA_GET_WATCHDOG) so that a protocol bug ends the kernel instead of hanging the GPU.
// A CTA's share of the pixels is a contiguous range of tiles.
#pragma once
#include <type_traits>

#include "gba_common.cuh"

#define GT_MAXDEG 21                // 6 * 21 + 1 = 127 rows <= 128
#ifndef GT_NPW
#define GT_NPW 14                   // producer warps 0 .. GT_NPW-1 (>= 13: the drain chunks sit on warps 0-12); 18 warps at
#endif                              // 96 registers measured 4 % slower, 20 spill
#define GT_ROT (GT_NPW == 14 ? 5 : 7) // slot-to-slot rotation of the edge assignment, coprime with GT_NPW
#define GT_LOADW GT_NPW             // loader warp
#define GT_MMAW (GT_NPW + 1)        // MMA-issue warp
#define GT_NT (32 * (GT_NPW + 2))
#define GT_K 32                     // pixels per operand stage (= producer lanes)
#define GT_STAGES 4                 // operand ring
#define GT_TPI 2                    // tiles per input stage
#define GT_IPX (GT_TPI * GT_K)      // pixels per input stage
#define GT_IS 3                     // input ring
#define GT_IN_ROWS (2 * GT_MAXDEG + 3)  // per edge the two weight rows (the E rows need no residual, hence no targets)
#define GT_IN_BYTES (GT_IN_ROWS * GT_IPX * 4)
#define GT_GROUP 4                  // tiles per accumulator group (16 truncating accumulations of hi hi^T)
#define GT_SBO 128
#define GT_LBO 2064
#define GT_BUF (8 * GT_LBO)         // one 128-row x 32-pixel tf32 operand tile
#define GT_STAGE_BYTES (2 * GT_BUF) // hi + lo
#define GT_TMEM_COLS 512

__host__ __device__ __forceinline__ size_t gram_tc_smem(int max_deg) {
  return (size_t)GT_STAGES * GT_STAGE_BYTES + (size_t)GT_IS * GT_IN_BYTES + 256 +
         (size_t)(max_deg < GT_MAXDEG ? max_deg : GT_MAXDEG) * (96 + sizeof(EdgeConst2)) + 32;
}

namespace tc {
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
// Ampere-style asynchronous 16-byte copy global -> shared (L2 only), and an mbarrier arrival that is
// deferred until all earlier cp.async of the executing thread have completed
__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
// try_wait with a suspend-time hint: the warp sleeps in hardware instead of burning issue slots
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(ns)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {  // the allocating warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle, sm_100 version bit
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor: D = f32, A = B = tf32, both K-major, M x N
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
// arrives on the mbarrier once every tcgen05 operation issued so far by this thread has completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 columns of fp32: thread t of the warp receives row (lane quarter base + t)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
}  // namespace tc

// A wait that is still unsatisfied after GT_WATCHDOG probes (>= 10 ms; a healthy wait is microseconds)
// means the pipeline protocol is broken.  Rather than hang the SM -- and with it the process and the
// GPU -- the warp records where it was and leaves; the others follow as their own waits expire, the
// CTA still frees its tensor memory, the iteration is treated like a failed factorisation (zero
// update, gba_info.chol_failures) and the host finds the records in GBA_GET_WATCHDOG.
#define GT_WATCHDOG 25000
__device__ __noinline__ void gram_tc_stuck(const BaParams &p, int site, int a, int b) {
  if ((threadIdx.x & 31) == 0) {
    *p.fail = 4;  // the iteration ends like a failed factorisation: zero pose update, counted in chol_failures
    const int n = atomicAdd(p.wdog, 1);
    if (n < 42) {
      int *r = p.wdog + 1 + 6 * n;
      r[0] = blockIdx.y; r[1] = blockIdx.x; r[2] = threadIdx.x >> 5; r[3] = site; r[4] = a; r[5] = b;
    }
  }
}

template <int MODE, bool IL>
__global__ void __launch_bounds__(GT_NT, 1) k_gram_tc(BaParams p) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  const int k = blockIdx.y;
  if (k >= p.st->K) return;
  if (p.owned && !p.owned[k]) return;
  const int e0 = p.eptr[k], deg = p.eptr[k + 1] - e0;
  if (deg == 0 || deg > GT_MAXDEG) return;  // larger keyframes are k_gram's
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int Rw = 6 * deg, R = Rw + 1;
  const int Rp = ((R + 7) / 8) * 8;          // pitch of G in the gram buffer (k_prep_scan)
  const int N = ((R + 15) / 16) * 16;        // MMA N: multiple of 16 for M = 128
  const int HW = p.HW, frame = p.kx[k];
  const int ntile = (HW + GT_K - 1) / GT_K;
  // this CTA's contiguous range of tiles
  const int tile0 = (int)(((long long)blockIdx.x * ntile) / gridDim.x);
  const int nloc = (int)(((long long)(blockIdx.x + 1) * ntile) / gridDim.x) - tile0;
  if (nloc <= 0) return;
  const int ngroup = (nloc + GT_GROUP - 1) / GT_GROUP;
  float *G = p.gram + p.goff[k];

  unsigned char *s_ops = s_raw;                                         // [GT_STAGES][hi | lo]
  float *s_in = reinterpret_cast<float *>(s_raw + (size_t)GT_STAGES * GT_STAGE_BYTES);  // [GT_IS][rows][GT_IPX]
  // barriers: full[S], empty[S], accfull[2], accfree[2], in_full[IS], in_empty[IS]
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_raw + (size_t)GT_STAGES * GT_STAGE_BYTES + (size_t)GT_IS * GT_IN_BYTES);
  uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_bar + 2 * GT_STAGES + 4 + 2 * GT_IS);
  EdgeGeo *s_geo = reinterpret_cast<EdgeGeo *>(reinterpret_cast<unsigned char *>(s_bar) + 256);
  EdgeConst2 *s_ec2 = reinterpret_cast<EdgeConst2 *>(s_geo + deg);  // per edge: R, t duplicated into f32x2 operands
  const float **s_row = reinterpret_cast<const float **>(s_ec2 + deg);  // [2 deg + 3] input row bases
  const uint32_t bar0 = tc::smem_u32(s_bar);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (GT_STAGES + s); };
  auto accfull_bar = [&](int a) { return bar0 + 8u * (2 * GT_STAGES + a); };
  auto accfree_bar = [&](int a) { return bar0 + 8u * (2 * GT_STAGES + 2 + a); };
  auto infull_bar = [&](int s) { return bar0 + 8u * (2 * GT_STAGES + 4 + s); };
  auto inempty_bar = [&](int s) { return bar0 + 8u * (2 * GT_STAGES + 4 + GT_IS + s); };

  if (warp == GT_MMAW) {
    if (lane == 0) {
      for (int s = 0; s < GT_STAGES; s++) {
        tc::mbar_init(full_bar(s), (uint32_t)min(deg, GT_NPW));  // one arrival per producer warp that has an edge
        tc::mbar_init(empty_bar(s), 1u);
      }
      for (int a = 0; a < 2; a++) {
        tc::mbar_init(accfull_bar(a), 1u);
        tc::mbar_init(accfree_bar(a), (uint32_t)GT_NPW);
      }
      for (int s = 0; s < GT_IS; s++) {
        tc::mbar_init(infull_bar(s), 32u);                         // the loader's 32 lanes, each after its copies landed
        tc::mbar_init(inempty_bar(s), (uint32_t)GT_NPW);           // every producer warp, edge or not
      }
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc<GT_TMEM_COLS>(tc::smem_u32(s_tmem));
  }
  // rows R .. Rp-1 (G's padding) are never produced: keep them zero in every stage
  for (int i = tid; i < (Rp - R) * GT_K * 2 * GT_STAGES; i += GT_NT) {
    const int r = R + i % (Rp - R), kk = (i / (Rp - R)) % GT_K, b = i / ((Rp - R) * GT_K);
    *reinterpret_cast<float *>(s_ops + (size_t)b * GT_BUF + (r & 7) * 16 + (r >> 3) * GT_SBO + (kk & 3) * 4 + (kk >> 2) * GT_LBO) = 0.f;
  }
  tc::fence_async_smem();
  for (int c = tid; c < 2 * deg + 3; c += GT_NT) {
    const float *src;
    if (c < 2 * deg) {
      const int n = p.eidx[e0 + (c >> 1)], a = c & 1;
      // interleaved layout: rows 2q, 2q+1 of a slot are the two 64-float halves of ONE run of 128 floats
      // (u0 v0 u1 v1 ...) of the edge's weights
      src = IL ? p.weights + (size_t)n * 2 * HW + a * GT_IPX : p.weights + ((size_t)n * 2 + a) * HW;
    } else {
      const int a = c - 2 * deg;
      src = (a == 0 ? p.disps + (size_t)frame * HW : a == 1 ? p.Q + (size_t)k * HW : p.W + (size_t)k * HW);
    }
    s_row[c] = src;
  }
  for (int i = tid; i < deg * 16; i += GT_NT)
    reinterpret_cast<float *>(s_geo)[i] = reinterpret_cast<const float *>(p.geo + p.eidx[e0 + (i >> 4)])[i & 15];
  for (int i = tid; i < deg; i += GT_NT) {
    const EdgeGeo *gg = p.geo + p.eidx[e0 + i];
    s_ec2[i] = make_edge_const2(gg->R, gg->t);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *s_tmem;

  if (warp < GT_NPW) {
    // ---------------- producers: items g = it * deg + q, warp w takes g = w, w + 14, ... ------------
    const PixConst pk = make_pix_const(p.intr);
    const float cx = p.intr[2], cy = p.intr[3], inv_wd = 1.0f / (float)p.wd;
    // this warp's corner of the accumulators: TMEM lanes 32 lq .., columns 32 cb ..
    const int lq = warp & 3, cb = lq + (warp >> 2);
    const bool has_chunk = cb < 4 && 32 * cb < N;
    const uint32_t tchunk = tmem + ((uint32_t)(32 * lq) << 16) + (uint32_t)(32 * cb);
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0.f;
    int fg = 0;  // next accumulator group to drain
    auto drain = [&]() {
      const int set = fg & 1;
      tc::fence_after_sync();
      if (has_chunk) {
        float v[32];
        tc::tmem_ld32(tchunk + (uint32_t)(256 * set), v);  // D1: hi hi^T
#pragma unroll
        for (int i = 0; i < 32; i++) acc[i] += v[i];
        tc::tmem_ld32(tchunk + (uint32_t)(256 * set + 128), v);  // D2: hi lo^T + lo hi^T
#pragma unroll
        for (int i = 0; i < 32; i++) acc[i] += v[i];
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(accfree_bar(set));
      fg++;
    };
    // Barrier probes are made warp-uniform with a vote: lanes can observe a completing phase a probe
    // apart, and drain() holds warp-collective tcgen05.ld -- a diverged warp there hangs the SM.
    auto poll_drain = [&]() {
      if (fg < ngroup && __all_sync(0xffffffffu, tc::mbar_test(accfull_bar(fg & 1), (uint32_t)(fg >> 1) & 1u))) drain();
    };
    auto probe = [&](uint32_t bar, uint32_t par) { return __all_sync(0xffffffffu, tc::mbar_try(bar, par, 400u)); };
    // byte offset of (row 0, k = lane) inside an operand tile
    const uint32_t lane_off = (uint32_t)(lane & 3) * 4u + (uint32_t)(lane >> 2) * GT_LBO;
    // Per input slot (GT_TPI tiles) a warp rebuilds the rows of one or two edges: ONE round of barrier
    // probes, one proxy fence and one arrival per slot and warp, with up to 2 x GT_TPI independent
    // pixels per lane in between -- a warp is a serial chain of ~100-cycle probes, so the work per probe
    // decides the throughput.  Which warps get two edges rotates with the slot.
    auto fast_slot = [&](auto ne_c, int is, int it0, int b0) {
      constexpr int NE = decltype(ne_c)::value;
      const float *in = s_in + (size_t)is * (GT_IN_BYTES / 4) + lane;
      float cwu[NE][GT_TPI], cwv[NE][GT_TPI], czq[GT_TPI], cqv[GT_TPI], cwk[GT_TPI];
#pragma unroll
      for (int u = 0; u < GT_TPI; u++) {
#pragma unroll
        for (int h = 0; h < NE; h++) {
          const int q = b0 + h * GT_NPW;
          if (IL) {  // (u, v) pairs: pixel u * 32 + lane of the slot sits at float 2 * (u * 32 + lane) of the run
            const float2 ww = *reinterpret_cast<const float2 *>(in + lane + (2 * q) * GT_IPX + 2 * u * GT_K);
            cwu[h][u] = ww.x; cwv[h][u] = ww.y;
          } else {
            cwu[h][u] = in[(2 * q + 0) * GT_IPX + u * GT_K];
            cwv[h][u] = in[(2 * q + 1) * GT_IPX + u * GT_K];
          }
        }
        czq[u] = in[(2 * deg + 0) * GT_IPX + u * GT_K];
        cqv[u] = in[(2 * deg + 1) * GT_IPX + u * GT_K];
        cwk[u] = (b0 == 0) ? in[(2 * deg + 2) * GT_IPX + u * GT_K] : 0.f;
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(inempty_bar(is));
      // per edge: row offsets inside an operand tile and the validity gate (degenerate edges give zero rows)
      uint32_t off[NE][6];
      float gate[NE];
#pragma unroll
      for (int h = 0; h < NE; h++) {
        const int q = b0 + h * GT_NPW;
        gate[h] = (s_geo[q].flags & 1) ? 0.f : 1.f;
#pragma unroll
        for (int c = 0; c < 6; c++) {
          const int r = 6 * q + c;
          off[h][c] = (uint32_t)(r & 7) * 16u + (uint32_t)(r >> 3) * GT_SBO;
        }
      }
      // the two pixels of the slot (tiles it0, it0 + 1) go through the arithmetic as one f32x2 pair
      static_assert(GT_TPI == 2, "the producers pair the two tiles of a slot");
      float csq[2], xs[2], ys[2];
      unsigned char *hi[2], *lo[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int it = it0 + u, stage = it % GT_STAGES;
        const int pix = (tile0 + it) * GT_K + lane;
        csq[u] = sqrtf(cqv[u]);
        const int prow = (int)(((float)pix + 0.5f) * inv_wd), pcol = pix - prow * p.wd;  // exact for pix < 2^22
        xs[u] = ((float)pcol - cx) * pk.ifx;
        ys[u] = ((float)prow - cy) * pk.ify;
        hi[u] = s_ops + (size_t)stage * GT_STAGE_BYTES + lane_off;
        lo[u] = hi[u] + GT_BUF;
      }
      const unsigned long long x2 = pack2(xs[0], xs[1]), y2 = pack2(ys[0], ys[1]), z2 = pack2(czq[0], czq[1]);
#pragma unroll
      for (int h = 0; h < NE; h++) {
        const EdgeConst2 &ec = s_ec2[b0 + h * GT_NPW];
        PixHat2 L;
        pix_hat2<MODE>(ec, pk, x2, y2, z2, pack2(cwu[h][0], cwu[h][1]), pack2(cwv[h][0], cwv[h][1]), L);
        const unsigned long long g2 = pack2(csq[0] * gate[h], csq[1] * gate[h]);
        const unsigned long long su = mul2(mul2(g2, L.wu), L.jzu), sv = mul2(mul2(g2, L.wv), L.jzv);
        const unsigned long long m1 = dup2(-1.0f), nsu = mul2(su, m1), nsv = mul2(sv, m1);
        // rows of sqrt(Q) E: su ju' + sv jv' with ju' = (e, 0, -ae, -ab, opa, -b), jv' = (0, e, -be, -opb, ab, a)
        unsigned long long e2[6];
        e2[0] = mul2(su, L.e);
        e2[1] = mul2(sv, L.e);
        e2[2] = fma2r(nsu, L.ae, mul2(nsv, L.be));
        e2[3] = fma2r(nsu, L.ab, mul2(nsv, L.opb));
        e2[4] = fma2r(su, L.opa, mul2(sv, L.ab));
        e2[5] = fma2r(nsu, L.b, mul2(sv, L.a));
#pragma unroll
        for (int c = 0; c < 6; c++) {
          float ev[2];
          unpack2(e2[c], ev[0], ev[1]);
#pragma unroll
          for (int u = 0; u < 2; u++) {
            const float hv = tc::tf32_rna(ev[u]);
            *reinterpret_cast<float *>(hi[u] + off[h][c]) = hv;
            *reinterpret_cast<float *>(lo[u] + off[h][c]) = ev[u] - hv;
          }
        }
      }
      if (b0 == 0) {  // the extra row sqrt(Q) w_k rides with edge 0
        const uint32_t offw = (uint32_t)(Rw & 7) * 16u + (uint32_t)(Rw >> 3) * GT_SBO;
#pragma unroll
        for (int u = 0; u < 2; u++) {
          const float x = csq[u] * cwk[u], hv = tc::tf32_rna(x);
          *reinterpret_cast<float *>(hi[u] + offw) = hv;
          *reinterpret_cast<float *>(lo[u] + offw) = x - hv;
        }
      }
    };
    const int nslot = (nloc + GT_TPI - 1) / GT_TPI;
    for (int j = 0; j < nslot; j++) {
      const int b0 = (warp + GT_ROT * j) % GT_NPW;  // coprime rotation: the heavy slots of a warp are spread out
      const int is = j % GT_IS;
      const int it0 = j * GT_TPI, nt = min(GT_TPI, nloc - it0);
      // Every warp takes part in the input ring of every slot, also when it has no edge in it
      // (deg < 14): a parity wait only means something to a warp that is less than one ring
      // revolution behind the barrier, so the loader must not be able to lap anybody.
      {
        const uint32_t par = (uint32_t)(j / GT_IS) & 1u;
        for (int spins = 0; !probe(infull_bar(is), par); poll_drain())
          if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 1, j, fg); goto bail; }
      }
      // ... and the same for the operand ring: a warp that skipped the wait of tile it - 4 could take an
      // 'empty' barrier that is still two phases back for one that has completed.
#pragma unroll
      for (int u = 0; u < GT_TPI; u++) {  // the MMAs that read these operand stages must be done
        const int it = it0 + u;
        if (u < nt && it >= GT_STAGES) {
          const uint32_t par = (uint32_t)(it / GT_STAGES - 1) & 1u;
          for (int spins = 0; !probe(empty_bar(it % GT_STAGES), par); poll_drain())
            if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 2, it, fg); goto bail; }
        }
      }
      if (b0 >= deg) {
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(inempty_bar(is));
        poll_drain();
        continue;
      }
      const int ne = (b0 + GT_NPW < deg) ? 2 : 1;
      // Fast path (all but the last slot of an image whose pixel count is not a multiple of 64): both
      // tiles present and inside the image -- no per-pixel validity tests, no divergence, one or two
      // edges fixed at compile time.  ~20 % fewer issue slots than the general path below.
      if (nt == GT_TPI && (tile0 + it0 + GT_TPI) * GT_K <= HW) {
        if (ne == 2) fast_slot(std::integral_constant<int, 2>{}, is, it0, b0);
        else fast_slot(std::integral_constant<int, 1>{}, is, it0, b0);
      } else {
        const int ne = (b0 + GT_NPW < deg) ? 2 : 1;
        const float *in = s_in + (size_t)is * (GT_IN_BYTES / 4) + lane;
        float cwu[2][GT_TPI], cwv[2][GT_TPI], czq[GT_TPI], cqv[GT_TPI], cwk[GT_TPI];
  #pragma unroll
        for (int u = 0; u < GT_TPI; u++) {
  #pragma unroll
          for (int h = 0; h < 2; h++) {
            const int q = (h < ne) ? b0 + h * GT_NPW : b0;
            if (IL) {
              const float2 ww = *reinterpret_cast<const float2 *>(in + lane + (2 * q) * GT_IPX + 2 * u * GT_K);
              cwu[h][u] = ww.x; cwv[h][u] = ww.y;
            } else {
              cwu[h][u] = in[(2 * q + 0) * GT_IPX + u * GT_K];
              cwv[h][u] = in[(2 * q + 1) * GT_IPX + u * GT_K];
            }
          }
          czq[u] = in[(2 * deg + 0) * GT_IPX + u * GT_K];
          cqv[u] = in[(2 * deg + 1) * GT_IPX + u * GT_K];
          cwk[u] = (b0 == 0) ? in[(2 * deg + 2) * GT_IPX + u * GT_K] : 0.f;
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(inempty_bar(is));
  #pragma unroll
        for (int u = 0; u < GT_TPI; u++) {
          if (u < nt) {
            const int it = it0 + u, stage = it % GT_STAGES;
            const int pix = (tile0 + it) * GT_K + lane;
            const bool pv = pix < HW;
            const float csq = pv ? sqrtf(cqv[u]) : 0.f;
            // row = pix / wd without an integer division (exact for pix < 2^22)
            const int prow = (int)(((float)pix + 0.5f) * inv_wd), pcol = pix - prow * p.wd;
            const float x0 = ((float)pcol - cx) * pk.ifx, y0 = ((float)prow - cy) * pk.ify;
            unsigned char *hi = s_ops + (size_t)stage * GT_STAGE_BYTES + lane_off, *lo = hi + GT_BUF;
  #pragma unroll
            for (int h = 0; h < 2; h++) {
              if (h < ne) {
                const int q = b0 + h * GT_NPW;
                const EdgeGeo &ge = s_geo[q];
                float e[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (pv && !(ge.flags & 1)) {
                  PixHat L;
                  pix_hat<MODE>(ge.R, ge.t, pk, x0, y0, czq[u], 0.f, 0.f, cwu[h][u], cwv[h][u], L);  // residuals unused
                  const float su = csq * L.wu * L.jzu, sv = csq * L.wv * L.jzv;
                  e[0] = su * L.e;
                  e[1] = sv * L.e;
                  e[2] = fmaf(su, L.nae, sv * L.nbe);
                  e[3] = fmaf(su, L.nab, sv * L.nopb);
                  e[4] = fmaf(su, L.opa, sv * L.ab);
                  e[5] = fmaf(su, L.nb, sv * L.a);
                }
  #pragma unroll
                for (int c = 0; c < 6; c++) {
                  const int r = 6 * q + c;
                  const uint32_t off = (uint32_t)(r & 7) * 16u + (uint32_t)(r >> 3) * GT_SBO;
                  const float hv = tc::tf32_rna(e[c]);
                  *reinterpret_cast<float *>(hi + off) = hv;
                  *reinterpret_cast<float *>(lo + off) = e[c] - hv;
                }
              }
            }
            if (b0 == 0) {  // the extra row sqrt(Q) w_k rides with edge 0
              const uint32_t off = (uint32_t)(Rw & 7) * 16u + (uint32_t)(Rw >> 3) * GT_SBO;
              const float x = pv ? csq * cwk[u] : 0.f, hv = tc::tf32_rna(x);
              *reinterpret_cast<float *>(hi + off) = hv;
              *reinterpret_cast<float *>(lo + off) = x - hv;
            }
          }
        }
      }
      tc::fence_async_smem();
      __syncwarp();
      if (lane == 0) {
#pragma unroll
        for (int u = 0; u < GT_TPI; u++)
          if (u < nt) tc::mbar_arrive(full_bar((it0 + u) % GT_STAGES));
      }
      poll_drain();
    }
    while (fg < ngroup) {  // the remaining groups, in order
      for (int spins = 0; !probe(accfull_bar(fg & 1), (uint32_t)(fg >> 1) & 1u);)
        if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 3, fg, ngroup); goto bail; }
      drain();
    }
    // ---------------- registers -> G (block-upper triangle; the mirror image is never read) ---------
    const int r = 32 * lq + lane;
    if (has_chunk && r < R) {
      const bool split = gridDim.x > 1;
      float *dst = G + (size_t)r * Rp + 32 * cb;
#pragma unroll
      for (int c = 0; c < 32; c += 4) {
        if (32 * cb + c < Rp) {  // Rp is a multiple of 8
          if (split) {
            atomicAdd(dst + c, acc[c]);
            atomicAdd(dst + c + 1, acc[c + 1]);
            atomicAdd(dst + c + 2, acc[c + 2]);
            atomicAdd(dst + c + 3, acc[c + 3]);
          } else {
            *reinterpret_cast<float4 *>(dst + c) = make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
          }
        }
      }
    }
  } else if (warp == GT_LOADW) {
    // ---------------- loader: bulk copies of the raw inputs, GT_IPX pixels per slot -----------------
    // Lane l moves 16 bytes: two 64-pixel rows per warp instruction.  (A TMA bulk copy per row was
    // tried first: at 256 B per copy the loader spent ~90 cycles per copy and set the kernel's pace.)
    const int nin = (nloc + GT_TPI - 1) / GT_TPI, nrow = 2 * deg + 3;
    const int sub = lane >> 4, c4 = (lane & 15) * 4;  // which row of the pair, first pixel of the 16 B piece
    for (int j = 0; j < nin; j++) {
      const int is = j % GT_IS, use = j / GT_IS;
      if (use > 0)
        for (int spins = 0; !__all_sync(0xffffffffu, tc::mbar_try(inempty_bar(is), (uint32_t)(use - 1) & 1u, 400u));)
          if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 4, j, 0); goto bail; }
      const int start = (tile0 + j * GT_TPI) * GT_K;
      const int npx = min(min(GT_IPX, HW - start), (nloc - j * GT_TPI) * GT_K);  // HW % 4 == 0: whole 16 B pieces
      const uint32_t dst0 = tc::smem_u32(s_in) + (uint32_t)is * GT_IN_BYTES + (uint32_t)c4 * 4u;
      // edge rows (weights), then the three keyframe rows.  Both loops are straight-line
      // (no test inside) so that the row-pointer loads of the unrolled iterations overlap: this warp is
      // a latency chain, a branch or a predicate in the body costs the whole kernel 15-60 %.
      const int nedge = 2 * deg;
      const int estart = IL ? 2 * start : start;  // interleaved: an edge row advances 2 floats per pixel
      const int elim = IL ? 2 * npx : npx;
      // c = sub, sub + 2, ...: the parity of c is this lane's `sub`, so the bound test leaves the loop
      const int eoff = IL ? sub * GT_IPX + c4 : c4;  // interleaved: rows 2q, 2q+1 are halves of one run
      if (eoff < elim) {
#pragma unroll 8
        for (int c = sub; c < nedge; c += 2) tc::cp_async16(dst0 + (uint32_t)c * (GT_IPX * 4), s_row[c] + estart + c4);
      }
      if (c4 < npx)
        for (int c = nedge + sub; c < nrow; c += 2) tc::cp_async16(dst0 + (uint32_t)c * (GT_IPX * 4), s_row[c] + start + c4);
      tc::cp_async_arrive(infull_bar(is));  // fires when this lane's copies above have landed
    }
  } else {
    // ---------------- MMA issue: one thread; per stage 4 k-steps x (D1 += hi hi^T, D2 += hi lo^T + lo hi^T) --
    const uint32_t idesc = tc::idesc_tf32(128, N);
    const uint32_t ops0 = tc::smem_u32(s_ops);
    for (int it = 0; it < nloc; it++) {
      const int stage = it % GT_STAGES, use = it / GT_STAGES;
      const int grp = it / GT_GROUP, set = grp & 1, first = (it % GT_GROUP) == 0;
      if (first && grp >= 2)  // the set must have been drained by every producer warp
        for (int spins = 0; !__all_sync(0xffffffffu, tc::mbar_try(accfree_bar(set), (uint32_t)((grp >> 1) - 1) & 1u, 400u));)
          if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 5, it, grp); goto bail; }
      for (int spins = 0; !__all_sync(0xffffffffu, tc::mbar_try(full_bar(stage), (uint32_t)use & 1u, 400u));)
        if (++spins > GT_WATCHDOG) { gram_tc_stuck(p, 6, it, grp); goto bail; }
      tc::fence_after_sync();
      if (lane == 0) {
        const uint32_t hi = ops0 + (uint32_t)stage * GT_STAGE_BYTES, lo = hi + GT_BUF;
        const uint32_t d1 = tmem + (uint32_t)(256 * set), d2 = d1 + 128u;
#pragma unroll
        for (int j = 0; j < GT_K / 8; j++) {
          const uint64_t dh = tc::smem_desc(hi + (uint32_t)j * 2u * GT_LBO, GT_LBO, GT_SBO);
          const uint64_t dl = tc::smem_desc(lo + (uint32_t)j * 2u * GT_LBO, GT_LBO, GT_SBO);
          const uint32_t accum = !(first && j == 0);
          tc::mma_tf32(d1, dh, dh, idesc, accum);
          tc::mma_tf32(d2, dh, dl, idesc, accum);
          tc::mma_tf32(d2, dl, dh, idesc, 1u);
        }
        tc::mma_commit(empty_bar(stage));
        if ((it % GT_GROUP) == GT_GROUP - 1 || it == nloc - 1) tc::mma_commit(accfull_bar(set));
      }
      __syncwarp();
    }
  }
bail:
  tc::fence_before_sync();
  __syncthreads();
  if (warp == GT_MMAW) tc::tmem_dealloc<GT_TMEM_COLS>(tmem);
}
