// gba_api.cu -- host side of libgba.so: context, launch sequencing, the extern "C" boundary
// declared in include/gba.h.  No torch types, no CPU compute path.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/gba.h"
#include "gba_common.cuh"
#include "gba_dspo.cuh"
#include "gba_kernels.cuh"
#include "gba_gram_tc.cuh"
#include "gba_solver.cuh"

thread_local char g_create_err[512] = "";  // also written by gba_noalloc.cu

struct gba_ctx {
  int device, N_max, E_max, ht, wd, HW, sm_count;
  // index structures
  int *kx, *fmap, *deg, *eptr, *eidx, *ek, *cc_list;
  unsigned long long *goff;
  Status *st_d, *st_h;
  // workspaces
  EdgeGeo *geo;
  float *hpart, *edge_hv, *Q, *W, *gram, *dx, *dzbuf, *dspo_part, *dwq;
  size_t gram_cap, sys_cap;
  double *sys;
  int *tslot, *perm, *iperm, *dspo_bad, *wdog;
  int *h_cnt;  // pinned: colcnt[nT_max] | rowcnt[nT_max], read back with the status
  int pos_cap;
  SolverBufs sb;
  unsigned int *flags;  // flagA | cntB | cntD | flagX | ticket, nT_max each
  unsigned int *upd;    // per tile slot: slab updates received inside k_chain
  size_t slot_cap;
  int nT_max, hpart_chunks, panel_capacity, chain_capacity;
  // state of the current call
  BaParams p;
  cudaStream_t stream;
  cudaStream_t side;           // k_gram (the few high-degree keyframes) runs here, beside k_gram_tc
  cudaEvent_t ev_fork, ev_join;
  int begun, launches, variant, gsplit, total_fail, last_P;
  int seg_lo, seg_n, first_rank;  // edge-sharded solve: this rank's band segments (seg_n = 0: not sharded)
  size_t sys_count;
  char err[512];
  // optional stage timing
  int timing, cur_stage, cur_launch0;
  std::vector<cudaEvent_t> *ev_pool;
  std::vector<int> *ev_stage;   // stage of event pair i (events 2i, 2i+1)
  std::vector<int> *ev_launch;  // launches inside pair i
  size_t ev_used;
};

// RAII-free stage brackets: stage_begin/stage_end record an event pair on the call's stream
static void stage_begin(gba_ctx *c, int stage) {
  if (!c->timing) return;
  if (c->ev_used + 2 > c->ev_pool->size()) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    c->ev_pool->push_back(a);
    c->ev_pool->push_back(b);
  }
  cudaEventRecord((*c->ev_pool)[c->ev_used], c->stream);
  c->cur_stage = stage;
  c->cur_launch0 = c->launches;
}
static void stage_end(gba_ctx *c) {
  if (!c->timing) return;
  cudaEventRecord((*c->ev_pool)[c->ev_used + 1], c->stream);
  c->ev_stage->push_back(c->cur_stage);
  c->ev_launch->push_back(c->launches - c->cur_launch0);
  c->ev_used += 2;
}

static int fail_msg(gba_ctx *c, int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(c ? c->err : g_create_err, 512, fmt, ap);
  va_end(ap);
  return code;
}

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail_msg(c, GBA_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

#define LAUNCH_CHECK(name)                                                                         \
  do {                                                                                             \
    c->launches++;                                                                                 \
    cudaError_t e_ = cudaGetLastError();                                                           \
    if (e_ != cudaSuccess) return fail_msg(c, GBA_ECUDA, "launch of %s failed: %s", name, cudaGetErrorString(e_)); \
  } while (0)

template <typename T>
static cudaError_t dalloc(T **p, size_t n) {
  return cudaMalloc((void **)p, std::max<size_t>(n, 1) * sizeof(T));
}

static const int kSolverSmem = DM_SMEM_BYTES;
static const int kGramSmemMax = 200 * 1024;

extern "C" const char *gba_last_error(const gba_ctx *c) { return c ? c->err : g_create_err; }

extern "C" void gba_destroy(gba_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  void *ptrs[] = {c->cc_list, c->kx, c->fmap, c->deg, c->eptr, c->eidx, c->ek, c->goff, c->st_d, c->geo, c->hpart, c->edge_hv,
                  c->Q, c->W, c->gram, c->dx, c->dzbuf, c->dspo_part, c->dwq, c->sys, c->tslot, c->perm, c->iperm, c->dspo_bad, c->wdog,
                  c->sb.tbits, c->sb.colrows, c->sb.colcnt, c->sb.rowcols, c->sb.rowcnt, c->sb.invL, c->sb.fail,
                  c->flags, c->upd, c->sb.trace};
  for (void *q : ptrs)
    if (q) cudaFree(q);
  if (c->side) cudaStreamDestroy(c->side);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->st_h) cudaFreeHost(c->st_h);
  if (c->h_cnt) cudaFreeHost(c->h_cnt);
  if (c->ev_pool) {
    for (cudaEvent_t e : *c->ev_pool) cudaEventDestroy(e);
    delete c->ev_pool;
    delete c->ev_stage;
    delete c->ev_launch;
  }
  free(c);
}

extern "C" int gba_create(gba_ctx **out, int device, int max_frames, int max_edges, int ht, int wd) {
  gba_ctx *c = nullptr;
  if (!out) return fail_msg(c, GBA_EINVAL, "out is NULL");
  *out = nullptr;
  if (max_frames < 2 || max_edges < 0 || ht < 1 || wd < 1)
    return fail_msg(c, GBA_EINVAL, "bad sizes: frames %d edges %d grid %dx%d", max_frames, max_edges, ht, wd);
  if (wd % 4 != 0) return fail_msg(c, GBA_EINVAL, "wd must be a multiple of 4 (got %d)", wd);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    return fail_msg(c, GBA_ECUDA, "no usable CUDA device %d (found %d); libgba has no CPU path", device, ndev);
  }
  gba_ctx *ctx = (gba_ctx *)calloc(1, sizeof(gba_ctx));
  c = ctx;
  c->ev_pool = new std::vector<cudaEvent_t>();
  c->ev_stage = new std::vector<int>();
  c->ev_launch = new std::vector<int>();
  c->device = device;
  c->N_max = max_frames;
  c->E_max = max_edges;
  c->ht = ht;
  c->wd = wd;
  c->HW = ht * wd;
  const size_t N = max_frames, E = std::max(max_edges, 1), HW = c->HW;
  c->pos_cap = max_frames + 26 * ORD_ALIGN;  // room for the padding of the segments and of the second-level separators
  c->nT_max = (6 * c->pos_cap + GBA_NB - 1) / GBA_NB;
  const size_t nT = c->nT_max, ld = nT * GBA_NB;
  if (nT > TS_MAXT) {
    free(ctx);
    return fail_msg(nullptr, GBA_ECAPACITY, "max_frames %d too large (limit %d)", max_frames, TS_MAXT * GBA_NB / 6);
  }
#define CKC(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      fail_msg(nullptr, GBA_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_));                \
      gba_destroy(ctx);                                                                            \
      return GBA_ECUDA;                                                                            \
    }                                                                                              \
  } while (0)
  CKC(cudaSetDevice(device));
  CKC(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
  c->hpart_chunks = (int)((HW + 511) / 512);
  c->gram_cap = std::min<size_t>((size_t)(6 * 64 + 8) * (6 * E + 8 * N), (size_t)1 << 30);
  c->sys_cap = std::min<size_t>(nT * (nT + 1) / 2, (size_t)131072) * GBA_NB2 + ld;
  CKC(dalloc(&c->kx, N));
  CKC(dalloc(&c->fmap, N));
  CKC(dalloc(&c->deg, N));
  CKC(dalloc(&c->cc_list, N));
  CKC(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  CKC(dalloc(&c->eptr, N + 1));
  CKC(dalloc(&c->eidx, E));
  CKC(dalloc(&c->ek, E));
  CKC(dalloc(&c->goff, N));
  CKC(dalloc(&c->st_d, 1));
  CKC(cudaMallocHost((void **)&c->st_h, sizeof(Status)));
  CKC(cudaMallocHost((void **)&c->h_cnt, 2 * nT * sizeof(int)));
  CKC(dalloc(&c->geo, E));
  CKC(dalloc(&c->hpart, E * c->hpart_chunks * 32));
  CKC(dalloc(&c->edge_hv, E * 32));
  CKC(dalloc(&c->Q, N * HW));
  CKC(dalloc(&c->W, N * HW));
  CKC(dalloc(&c->gram, c->gram_cap));
  CKC(dalloc(&c->dx, 6 * N));
  CKC(dalloc(&c->dzbuf, N * HW));
  CKC(dalloc(&c->dspo_part, N * c->hpart_chunks * 8));
  CKC(dalloc(&c->dwq, 2 * N));
  CKC(dalloc(&c->sys, c->sys_cap));
  CKC(dalloc(&c->sb.invL, nT * GBA_NB2));
  CKC(dalloc(&c->tslot, nT * nT));
  CKC(dalloc(&c->perm, N));
  CKC(dalloc(&c->iperm, (size_t)c->pos_cap));
  CKC(dalloc(&c->sb.colrows, nT * nT));
  CKC(dalloc(&c->sb.colcnt, nT));
  CKC(dalloc(&c->sb.rowcols, nT * nT));
  CKC(dalloc(&c->sb.rowcnt, nT));
  CKC(dalloc(&c->sb.fail, 1));
  CKC(dalloc(&c->dspo_bad, 1));
  CKC(dalloc(&c->wdog, 256));
  CKC(cudaMemset(c->wdog, 0, 256 * sizeof(int)));
  CKC(dalloc(&c->sb.tbits, nT * TS_WORDS));
  CKC(dalloc(&c->flags, 7 * nT));
  c->slot_cap = std::min<size_t>(nT * (nT + 1) / 2, (size_t)131072);
  CKC(dalloc(&c->upd, c->slot_cap));
  c->sb.upd = c->upd;
  c->sb.trace = nullptr;
  if (getenv("GBA_CHAIN_TRACE")) {  // development aid: task timeline of k_chain (GBA_GET_CHAIN_TRACE)
    CKC(cudaMalloc((void **)&c->sb.trace, (1 + 4 * (size_t)CHAIN_TRACE_CAP) * sizeof(unsigned long long)));
    CKC(cudaMemset(c->sb.trace, 0, sizeof(unsigned long long)));
  }
  c->sb.ticket = c->flags + 4 * nT;
  c->sb.flagA = c->flags;
  c->sb.cntB = c->flags + nT;
  c->sb.cntD = c->flags + 2 * nT;
  c->sb.flagX = c->flags + 3 * nT;
  c->sb.cntY = c->flags + 5 * nT;
  c->sb.ticket2 = c->flags + 6 * nT;
  CKC(cudaMemset(c->sb.fail, 0, sizeof(int)));
  CKC(cudaMemset(c->st_d, 0, sizeof(Status)));
  CKC(cudaFuncSetAttribute(k_gram, cudaFuncAttributeMaxDynamicSharedMemorySize, kGramSmemMax));
  CKC(cudaFuncSetAttribute(k_gram_tc<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gram_tc_smem(GT_MAXDEG)));
  CKC(cudaFuncSetAttribute(k_gram_tc<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gram_tc_smem(GT_MAXDEG)));
  CKC(cudaFuncSetAttribute(k_gram_tc<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gram_tc_smem(GT_MAXDEG)));
  CKC(cudaFuncSetAttribute(k_gram_tc<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gram_tc_smem(GT_MAXDEG)));
  CKC(cudaFuncSetAttribute(k_chain<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSolverSmem));
  {
    int per_sm = 0;
    CKC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chain<16>, 256, kSolverSmem));
    c->chain_capacity = std::max(1, per_sm) * c->sm_count;
  }
#undef CKC
  *out = ctx;
  return GBA_OK;
}

// ---- index build shared by BA and DSPO ---------------------------------------------------------
// CSR of the edges in p.ii / p.jj by source keyframe
static int build_csr(gba_ctx *c, const int64_t *kx_global, int n_kx, bool first_pass = true) {
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  k_prep_frames<<<1, 1024, 0, s>>>(p, kx_global, n_kx, first_pass ? 1 : 0);
  LAUNCH_CHECK("k_prep_frames");
  if (p.E > 0) {
    k_prep_edges<<<(p.E + 255) / 256, 256, 0, s>>>(p);
    LAUNCH_CHECK("k_prep_edges");
  }
  k_prep_scan<<<1, 1024, 0, s>>>(p);
  LAUNCH_CHECK("k_prep_scan");
  if (p.E > 0) {
    k_prep_fill<<<(p.N * 32 + 255) / 256, 256, 0, s>>>(p);
    LAUNCH_CHECK("k_prep_fill");
  }
  return GBA_OK;
}

// ii_all/jj_all (optional): the whole graph, when p.ii/p.jj hold only this rank's shard; the pose
// ordering and the tile structure must be identical on all ranks, so they come from the whole graph
static int build_index(gba_ctx *c, const int64_t *kx_global, int n_kx, bool want_tiles,
                       const int64_t *ii_all = nullptr, const int64_t *jj_all = nullptr, int n_all = 0) {
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  stage_begin(c, GBA_STAGE_INDEX);
  const bool sharded = want_tiles && ii_all && jj_all;
  const int64_t *ii_loc = p.ii, *jj_loc = p.jj;
  const int E_loc = p.E;
  if (sharded) {
    p.ii = ii_all;
    p.jj = jj_all;
    p.E = n_all;
  }
  int rc = build_csr(c, kx_global, n_kx);
  if (rc) return rc;
  if (want_tiles) {
    k_order<<<1, 1024, (size_t)((p.P + 3) / 4) * 4, s>>>(p, c->pos_cap, getenv("GBA_NO_LEVEL2") ? 1 : 0);
    LAUNCH_CHECK("k_order");
    CK(cudaMemsetAsync(c->sb.tbits, 0, (size_t)p.nT * TS_WORDS * sizeof(unsigned int), s));
    k_tile_mark<false><<<p.N, 128, 0, s>>>(p, c->sb.tbits);
    LAUNCH_CHECK("k_tile_mark");
    k_tile_fill<<<1, 512, 0, s>>>(p, c->sb);
    LAUNCH_CHECK("k_tile_fill");
  }
  if (sharded) {  // now the CSR of this rank's own edges
    p.ii = ii_loc;
    p.jj = jj_loc;
    p.E = E_loc;
    rc = build_csr(c, kx_global, n_kx, false);
    if (rc) return rc;
    // every tile a local edge will add to must exist in the structure derived from the whole graph
    k_tile_mark<true><<<p.N, 128, 0, s>>>(p, c->sb.tbits);
    LAUNCH_CHECK("k_tile_mark<check>");
  }
  stage_end(c);
  CK(cudaMemcpyAsync(c->st_h, c->st_d, sizeof(Status), cudaMemcpyDeviceToHost, s));
  if (want_tiles) {
    CK(cudaMemcpyAsync(c->h_cnt, c->sb.colcnt, (size_t)c->nT_max * sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(c->h_cnt + c->nT_max, c->sb.rowcnt, (size_t)c->nT_max * sizeof(int), cudaMemcpyDeviceToHost, s));
  }
  CK(cudaStreamSynchronize(s));
  const Status &st = *c->st_h;
  if (st.err & 1) return fail_msg(c, GBA_EINVAL, "ii/jj (or kx) contain keyframe indices outside [0,%d)", p.N);
  if (st.err & 4) return fail_msg(c, GBA_EINVAL, "an edge's source keyframe is missing from kx_global");
  if (st.err & 8) return fail_msg(c, GBA_EINVAL, "a local edge is missing from ii_all / jj_all (the whole-graph edge list)");
  return GBA_OK;
}

extern "C" int gba_ba_begin(gba_ctx *c, float *poses, float *disps, const float *intrinsics, const float *disps_sens,
                            const float *targets, const float *weights, const float *eta, const int64_t *ii,
                            const int64_t *jj, int n_frames, int n_edges, int eta_rows, int t0, int t1, float lm,
                            float ep, int motion_only, int depth_only, int mode, const int64_t *kx_global, int n_kx,
                            const uint8_t *owned, const int64_t *ii_all, const int64_t *jj_all, int n_all,
                            void *stream) {
  if (!c) return GBA_EINVAL;
  c->begun = 0;
  if (!poses || !disps || !intrinsics || !disps_sens || !eta || (n_edges > 0 && (!targets || !weights || !ii || !jj)))
    return fail_msg(c, GBA_EINVAL, "NULL array argument");
  if (n_frames < 1 || n_frames > c->N_max)
    return fail_msg(c, GBA_ECAPACITY, "n_frames %d outside [1,%d]", n_frames, c->N_max);
  if (n_edges < 0 || n_edges > c->E_max) return fail_msg(c, GBA_ECAPACITY, "n_edges %d outside [0,%d]", n_edges, c->E_max);
  if (ii_all && jj_all && (n_all < n_edges || n_all > c->E_max))
    return fail_msg(c, GBA_ECAPACITY, "n_all %d outside [n_edges,%d]", n_all, c->E_max);
  if (t0 < 0 || t1 <= t0 || t1 > n_frames)
    return fail_msg(c, GBA_EINVAL, "optimised window [%d,%d) not inside [0,%d)", t0, t1, n_frames);
  if ((mode & ~GBA_LAYOUT_INTERLEAVED) != GBA_MODE_DROID_CUDA && (mode & ~GBA_LAYOUT_INTERLEAVED) != GBA_MODE_GEOM_PY)
    return fail_msg(c, GBA_EINVAL, "unknown mode %d", mode);
  // every per-pixel array is read with 16-byte vector loads (and 16-byte bulk copies in k_gram_tc)
  if ((((uintptr_t)disps | (uintptr_t)disps_sens | (uintptr_t)targets | (uintptr_t)weights | (uintptr_t)eta) & 15) != 0)
    return fail_msg(c, GBA_EINVAL, "disps, disps_sens, targets, weights and eta must be 16-byte aligned");
  if ((((uintptr_t)poses | (uintptr_t)intrinsics) & 3) != 0 || (((uintptr_t)ii | (uintptr_t)jj) & 7) != 0)
    return fail_msg(c, GBA_EINVAL, "misaligned poses / intrinsics / ii / jj pointer");
  CK(cudaSetDevice(c->device));
  c->stream = (cudaStream_t)stream;
  c->launches = 0;
  c->total_fail = 0;
  c->seg_lo = 0;
  c->seg_n = 0;
  c->first_rank = 1;
  BaParams &p = c->p;
  memset(&p, 0, sizeof(p));
  p.poses = poses; p.disps = disps; p.intr = intrinsics; p.sens = disps_sens; p.targets = targets;
  p.weights = weights; p.eta = eta; p.ii = ii; p.jj = jj;
  p.N = n_frames; p.E = n_edges; p.HW = c->HW; p.wd = c->wd; p.t0 = t0; p.t1 = t1; p.P = t1 - t0;
  p.mode = mode & ~GBA_LAYOUT_INTERLEAVED; p.il = (mode & GBA_LAYOUT_INTERLEAVED) ? 1 : 0;
  p.motion_only = motion_only ? 1 : 0; p.depth_only = depth_only ? 1 : 0; p.lm = lm; p.ep = ep;
  p.cc_list = c->cc_list; p.kx = c->kx; p.fmap = c->fmap; p.deg = c->deg; p.eptr = c->eptr; p.eidx = c->eidx; p.ek = c->ek;
  p.owned = owned; p.goff = c->goff; p.st = c->st_d; p.geo = c->geo; p.hpart = c->hpart; p.edge_hv = c->edge_hv;
  p.Q = c->Q; p.W = c->W; p.gram = c->gram; p.dx = c->dx; p.dz_out = c->dzbuf; p.sys = c->sys; p.tslot = c->tslot;
  p.perm = c->perm;
  p.iperm = c->iperm;
  p.nT = c->nT_max;  // stride of the tile tables; the actual tile count comes back in the status
  {  // GBA_GRAM_TC=0 sends every keyframe through the CUDA-core k_gram (diagnostics, A/B comparison)
    const char *e = getenv("GBA_GRAM_TC");
    p.tc_maxdeg = (e && e[0] == '0') ? 0 : GT_MAXDEG;
    p.wdog = c->wdog;
    p.fail = c->sb.fail;
  }
  p.ld = 0;
  c->last_P = p.P;
  int rc = build_index(c, kx_global, n_kx, true, ii_all, jj_all, n_all);
  if (rc) return rc;
  const Status &st = *c->st_h;
  if (!motion_only && eta_rows != st.K)
    return fail_msg(c, GBA_EINVAL, "eta has %d rows but the graph has K = %d depth keyframes", eta_rows, st.K);
  p.ld = st.n_tiles * GBA_NB;
  if ((size_t)st.n_slots * GBA_NB2 + p.ld > c->sys_cap)
    return fail_msg(c, GBA_ECAPACITY, "reduced system needs %d tiles, capacity exceeded", st.n_slots);
  c->sys_count = (size_t)st.n_slots * GBA_NB2 + p.ld;
  if (!motion_only) {
    if (st.gram_floats > c->gram_cap)
      return fail_msg(c, GBA_ECAPACITY, "Schur Gram buffers need %llu floats (capacity %zu); max out-degree %d",
                      st.gram_floats, c->gram_cap, st.max_deg);
    const int RS = gram_pitch((6 * st.max_deg + 1 + 7) / 8);
    if ((size_t)2 * GRAM_T * RS * sizeof(float) + (size_t)st.max_deg * 68 > (size_t)kGramSmemMax)
      return fail_msg(c, GBA_ECAPACITY, "out-degree %d too large for the shared-memory Gram tile", st.max_deg);
  }
  // launch shape of the streaming passes (measured at config E: 128 threads x 8 pixels beats 256 x 8 for the
  // linearisation, 0.39 vs 0.42 ms, and 128 x 4 is best for the back-substitution, 0.14 vs 0.19 ms):
  // k_linearize takes the 1024-pixel footprint when that still gives >= 2 CTAs per SM, else 512; k_backsub always 512
  const int K = std::max(st.K, 1), want = 2 * c->sm_count;
  const int foot[3] = {2048, 1024, 512};
  c->variant = ((long long)K * ((c->HW + 1023) / 1024) >= want) ? 1 : 2;
  if (const char *ev = getenv("GBA_VARIANT")) c->variant = std::max(0, std::min(2, atoi(ev)));  // diagnostics
  p.nchunk = (c->HW + foot[c->variant] - 1) / foot[c->variant];
  const int ntile = (c->HW + GRAM_T - 1) / GRAM_T;
  // Gram CTAs per keyframe: enough CTAs to fill the GPU (2 resident per SM), and among the small
  // factors the one with the least tail in the last wave of CTAs
  c->gsplit = std::max(1, std::min(ntile, (3 * c->sm_count + K - 1) / K));
  if (c->gsplit <= 4) {
    const int slots = 2 * c->sm_count;
    double best = 1e30;
    for (int g = 1; g <= std::min(4, ntile); g++) {
      const double waves = (double)((K * g + slots - 1) / slots) / g;
      if (waves < best - 0.05) {
        best = waves;
        c->gsplit = g;
      }
    }
  }
  c->begun = 1;
  return GBA_OK;
}

extern "C" int gba_ba_build(gba_ctx *c) {
  if (!c || !c->begun) return c ? fail_msg(c, GBA_EINVAL, "gba_ba_build without gba_ba_begin") : GBA_EINVAL;
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  const Status &st = *c->st_h;
  CK(cudaSetDevice(c->device));
  if (p.E > 0) {
    stage_begin(c, GBA_STAGE_POSE);
    k_edge_geo<<<(p.E + 255) / 256, 256, 0, s>>>(p);
    LAUNCH_CHECK("k_edge_geo");
    stage_end(c);
  }
  stage_begin(c, GBA_STAGE_LINEARIZE);  // brackets k_linearize alone (roofline kernel)
  if (st.K > 0) {
    dim3 grid(p.nchunk, st.K);
    // (semantics mode, layout) x (CTA footprint): compile-time variants, chosen once per launch
#define GBA_LAUNCH_K_LINEARIZE(MODE, IL)                                                    \
  do {                                                                                   \
    if (c->variant == 0) k_linearize<256, 2, MODE, IL><<<grid, 256, 0, s>>>(p);            \
    else if (c->variant == 1) k_linearize<128, 2, MODE, IL><<<grid, 128, 0, s>>>(p);       \
    else k_linearize<128, 1, MODE, IL><<<grid, 128, 0, s>>>(p);                            \
  } while (0)
    if (p.mode == 0 && !p.il) GBA_LAUNCH_K_LINEARIZE(0, false);
    else if (p.mode == 0) GBA_LAUNCH_K_LINEARIZE(0, true);
    else if (!p.il) GBA_LAUNCH_K_LINEARIZE(1, false);
    else GBA_LAUNCH_K_LINEARIZE(1, true);
#undef GBA_LAUNCH_K_LINEARIZE
    LAUNCH_CHECK("k_linearize");
  }
  stage_end(c);
  stage_begin(c, GBA_STAGE_POSE);
  CK(cudaMemsetAsync(c->sys, 0, c->sys_count * sizeof(double), s));
  k_pad_diag<<<(p.ld + 255) / 256, 256, 0, s>>>(p);
  LAUNCH_CHECK("k_pad_diag");
  if (p.E > 0) {
    k_edge_finish<<<(p.E + 127) / 128, 128, 0, s>>>(p);
    LAUNCH_CHECK("k_edge_finish");
  }
  if (p.mode == GBA_MODE_GEOM_PY) {  // damping on H before the Schur complement (chol.py:66-68)
    k_damp<<<(6 * p.P + 255) / 256, 256, 0, s>>>(p, p.lm, p.ep, 0, -1, 1, 0, 0);
    LAUNCH_CHECK("k_damp");
  }
  stage_end(c);
  if (!p.motion_only && p.E > 0 && st.K > 0) {
    stage_begin(c, GBA_STAGE_GRAM);
    // pixel split per keyframe.  k_gram_tc runs one CTA per SM and pays ~15 us of set-up per CTA, so it
    // is only split when the keyframes alone leave SMs idle; with it active, k_gram only sees the few
    // high-degree keyframes and is split wide.
    int gs_tc = 1, gs_cc = c->gsplit;
    if (p.tc_maxdeg > 0) {
      const int ntile = (c->HW + GT_K - 1) / GT_K;
      double best = 1e30;
      for (int g = 1; g <= std::min(16, ntile); g++) {
        const double cost = (double)((st.K * g + c->sm_count - 1) / c->sm_count) * (1.0 / g + 0.12);
        if (cost < best - 1e-9) {
          best = cost;
          gs_tc = g;
        }
      }
      gs_cc = std::max(c->gsplit, std::min(16, ntile));
    }
    // GBA_DETERMINISTIC=1 (parity runs): one CTA per keyframe in both Gram kernels, so that no fp32 sum is merged by
    // atomics and the fp32 inputs of the reduced system are reproducible run to run (slower when there are fewer
    // keyframes than SMs).  The remaining unordered sums are fp64 (pose blocks, Schur assembly, closure tiles, rhs).
    if (getenv("GBA_DETERMINISTIC")) gs_tc = gs_cc = 1;
    const bool need_cc = st.n_cc > 0;
    if (need_cc && !getenv("GBA_DETERMINISTIC"))  // few keyframes: split their pixels until the CTAs fill the machine
      gs_cc = std::min((c->HW + GRAM_T - 1) / GRAM_T, std::max(gs_cc, (2 * c->sm_count + st.n_cc - 1) / st.n_cc));
    if ((p.tc_maxdeg > 0 && gs_tc > 1) || (need_cc && gs_cc > 1))
      CK(cudaMemsetAsync(c->gram, 0, (size_t)st.gram_floats * sizeof(float), s));
    const bool both = need_cc && p.tc_maxdeg > 0;
    if (need_cc) {  // the keyframes with more out-edges than k_gram_tc takes (cc_list), beside it on the side stream
      cudaStream_t sc = both ? c->side : s;
      if (both) {
        CK(cudaEventRecord(c->ev_fork, s));
        CK(cudaStreamWaitEvent(sc, c->ev_fork, 0));
      }
      const int RS = gram_pitch((6 * st.max_deg + 1 + 7) / 8);
      k_gram<<<dim3(gs_cc, st.n_cc), GRAM_NT, (size_t)2 * GRAM_T * RS * sizeof(float) + (size_t)st.max_deg * 68, sc>>>(p);
      LAUNCH_CHECK("k_gram");
      if (both) CK(cudaEventRecord(c->ev_join, sc));
    }
    if (p.tc_maxdeg > 0) {  // tensor-core Gram for the keyframes with <= GT_MAXDEG out-edges
      const dim3 gg(gs_tc, st.K);
      const size_t sm = gram_tc_smem(st.max_deg);
      if (p.mode == GBA_MODE_DROID_CUDA && !p.il) k_gram_tc<0, false><<<gg, GT_NT, sm, s>>>(p);
      else if (p.mode == GBA_MODE_DROID_CUDA) k_gram_tc<0, true><<<gg, GT_NT, sm, s>>>(p);
      else if (!p.il) k_gram_tc<1, false><<<gg, GT_NT, sm, s>>>(p);
      else k_gram_tc<1, true><<<gg, GT_NT, sm, s>>>(p);
      LAUNCH_CHECK("k_gram_tc");
    }
    if (both) CK(cudaStreamWaitEvent(s, c->ev_join, 0));
    stage_end(c);
    stage_begin(c, GBA_STAGE_SCHUR);
    k_schur<<<st.K, SCHUR_NT, (size_t)(6 * (6 * st.max_deg + 1) + st.max_deg) * sizeof(float), s>>>(p);
    LAUNCH_CHECK("k_schur");
    stage_end(c);
  }
  return GBA_OK;
}

extern "C" int gba_ba_system(gba_ctx *c, double **buf, size_t *count) {
  if (!c || !c->begun || !buf || !count) return GBA_EINVAL;
  *buf = c->sys;
  *count = c->sys_count;
  return GBA_OK;
}

// The factorisation kernel.  phase 0: everything (one rank).  Edge-sharded runs: phase 1 = the tile columns of this
// rank's band segments, phase 2 = the closure block after the ranks have summed it.
static int launch_chain(gba_ctx *c, int phase) {
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  const Status &st = *c->st_h;
  const int nT = st.n_tiles, nseg = st.n_seg, segT = st.seg_tiles, band = nseg > 1 ? nseg * segT : 0;
  const int *colcnt = c->h_cnt;
  const int seg_lo = phase == 1 ? c->seg_lo : 0, seg_n = phase == 1 ? c->seg_n : (nseg > 1 ? nseg : 0);
  // spine queue: the first potrf of every chain, then per column its critical tasks; bulk queue: the rest
  long long total_spine = 0, total_bulk = 0;
  auto add_col = [&](int kk, bool has_next) {
    const int cnt = colcnt[kk];
    total_spine += (cnt > 0 ? 8 : 0) + (has_next ? 1 : 0);
    total_bulk += cnt > 0 ? 4LL * (cnt - 1) + (long long)cnt * (cnt + 1) / 2 - 1 : 0;
  };
  int chains = 0;
  if (phase != 2 && nseg > 1) {
    for (int sg = seg_lo; sg < seg_lo + seg_n; sg++)
      for (int t = 0; t < segT; t++) {
        const int kk = sg * segT + t;
        add_col(kk, t + 1 < segT || (phase == 0 && kk == band - 1 && band < nT));
      }
    chains = seg_n;
    total_spine += seg_n;
  }
  if (phase != 1) {
    for (int kk = band; kk < nT; kk++) add_col(kk, kk + 1 < nT);
    total_spine += st.n_seg2 + 1;  // (upper bounds: they only size the grid)
    if (chains == 0) chains = std::max(1, st.n_seg2);
  }
  if (total_bulk >= (1LL << 30)) return fail_msg(c, GBA_ECAPACITY, "reduced system too large for the tile task graph");
  CK(cudaMemsetAsync(c->flags, 0, (size_t)5 * c->nT_max * sizeof(unsigned int), s));  // flagA .. ticket (not cntY / ticket2)
  CK(cudaMemsetAsync(c->upd, 0, (size_t)st.n_slots * sizeof(unsigned int), s));
  if (c->sb.trace) CK(cudaMemsetAsync(c->sb.trace, 0, sizeof(unsigned long long), s));  // keeps the last launch
  // spine CTAs: 4 slabs in flight per independent chain (band segments run side by side), at least 8
  int spine_width = std::min(c->chain_capacity / 2, std::max(8, 4 * chains));
  if (const char *e = getenv("GBA_CHAIN_SPINE")) spine_width = std::min(c->chain_capacity / 2, std::max(2, atoi(e) * chains) & ~1);  // tuning aid
  const int n_bulk = (int)std::max<long long>(1, std::min<long long>(total_bulk, (long long)c->chain_capacity - spine_width));
  // a grid of exactly two CTAs on every SM: spine CTAs by SM (pairs) instead of by block index, see k_chain
  const bool by_sm = spine_width + n_bulk == 2 * c->sm_count && c->chain_capacity == 2 * c->sm_count &&
                     spine_width % 2 == 0 && !getenv("GBA_CHAIN_SPINE_BY_BLOCK");
  k_chain<16><<<spine_width + n_bulk, 256, kSolverSmem, s>>>(p, c->sb, nseg, segT, (int)total_spine, (int)total_bulk,
                                                           by_sm ? -spine_width : spine_width, phase, seg_lo,
                                                           phase == 1 ? c->seg_n : 0, st.n_seg2, st.seg2_tiles);
  LAUNCH_CHECK("k_chain");
  return GBA_OK;
}

// backward substitution (one launch), dx, failure accounting, then back-substitution of the disparities and the
// pose retraction
static int solve_tail(gba_ctx *c, bool sharded) {
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  const Status &st = *c->st_h;
  const int nT = st.n_tiles;
  const int *colcnt = c->h_cnt, *rowcnt = c->h_cnt + c->nT_max;
  stage_begin(c, GBA_STAGE_SOLVE);
  CK(cudaMemsetAsync(c->sb.cntY, 0, (size_t)2 * c->nT_max * sizeof(unsigned int), s));  // cntY, ticket2
  CK(cudaMemsetAsync(c->sb.flagX, 0, (size_t)c->nT_max * sizeof(unsigned int), s));
  {  // one launch: a task per active tile below the diagonal, plus the rows that receive no update
    int tasks = 0;
    for (int i = 0; i < nT; i++) tasks += rowcnt[i] + (colcnt[i] == 0 ? 1 : 0);
    const int grid = std::max(1, std::min(tasks, 4 * c->sm_count));
    k_back<<<grid, 256, 0, s>>>(p, c->sb, tasks, sharded ? c->seg_lo : 0, sharded ? c->seg_n : -1, st.n_seg, st.seg_tiles);
    LAUNCH_CHECK("k_back");
  }
  k_dx_out<<<(6 * p.P + 255) / 256, 256, 0, s>>>(p, c->sb, sharded ? c->seg_lo : 0, sharded ? c->seg_n : -1, st.n_seg, st.seg_tiles);
  LAUNCH_CHECK("k_dx_out");
  k_fail_account<<<1, 1, 0, s>>>(p, c->sb);
  LAUNCH_CHECK("k_fail_account");
  stage_end(c);
  if (!p.motion_only && st.K > 0) {
    stage_begin(c, GBA_STAGE_BACKSUB);
    dim3 grid((c->HW + 511) / 512, st.K);
    // (semantics mode, layout): compile-time variants, chosen once per launch
#define GBA_LAUNCH_K_BACKSUB(MODE, IL) k_backsub<128, 1, MODE, IL><<<grid, 128, 0, s>>>(p)
    if (p.mode == 0 && !p.il) GBA_LAUNCH_K_BACKSUB(0, false);
    else if (p.mode == 0) GBA_LAUNCH_K_BACKSUB(0, true);
    else if (!p.il) GBA_LAUNCH_K_BACKSUB(1, false);
    else GBA_LAUNCH_K_BACKSUB(1, true);
#undef GBA_LAUNCH_K_BACKSUB
    LAUNCH_CHECK("k_backsub");
    if (p.mode == GBA_MODE_GEOM_PY && st.K < p.N) {  // ba.py:119 clamps every row, k_backsub the rows in kx
      k_clamp_rest<<<dim3((c->HW / 4 + 127) / 128, p.N), 128, 0, s>>>(p.disps, c->fmap, c->HW);
      LAUNCH_CHECK("k_clamp_rest");
    }
    stage_end(c);
  }
  if (p.motion_only || !p.depth_only) {
    stage_begin(c, GBA_STAGE_RETR);
    k_pose_retr<<<(p.P + 127) / 128, 128, 0, s>>>(p);
    LAUNCH_CHECK("k_pose_retr");
    stage_end(c);
  }
  return GBA_OK;
}

extern "C" int gba_ba_solve(gba_ctx *c) {
  if (!c || !c->begun) return c ? fail_msg(c, GBA_EINVAL, "gba_ba_solve without gba_ba_begin") : GBA_EINVAL;
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  CK(cudaSetDevice(c->device));
  stage_begin(c, GBA_STAGE_FACTOR);
  if (p.mode == GBA_MODE_DROID_CUDA) {  // damping after the Schur complement (droid_kernels.cu:1197,1407)
    k_damp<<<(6 * p.P + 255) / 256, 256, 0, s>>>(p, p.lm, p.ep, 0, -1, 1, 0, 0);
    LAUNCH_CHECK("k_damp");
  }
  int rc = launch_chain(c, 0);
  if (rc) return rc;
  stage_end(c);
  return solve_tail(c, false);
}

// ---- the solve of an edge-sharded run, split at the one exchange it needs ------------------------------------
extern "C" int gba_ba_set_shard(gba_ctx *c, int seg_lo, int seg_n, int first_rank) {
  if (!c || !c->begun) return c ? fail_msg(c, GBA_EINVAL, "gba_ba_set_shard without gba_ba_begin") : GBA_EINVAL;
  const Status &st = *c->st_h;
  if (st.n_seg < 2) return fail_msg(c, GBA_EINVAL, "the reduced system of this graph has no independent band segments");
  if (seg_lo < 0 || seg_n < 1 || seg_lo + seg_n > st.n_seg)
    return fail_msg(c, GBA_EINVAL, "segments [%d,%d) outside [0,%d)", seg_lo, seg_lo + seg_n, st.n_seg);
  if (c->p.mode != GBA_MODE_DROID_CUDA) return fail_msg(c, GBA_EINVAL, "sharded solve: droid_cuda semantics only");
  c->seg_lo = seg_lo;
  c->seg_n = seg_n;
  c->first_rank = first_rank ? 1 : 0;
  return GBA_OK;
}

extern "C" int gba_ba_eliminate(gba_ctx *c) {
  if (!c || !c->begun || c->seg_n < 1) return c ? fail_msg(c, GBA_EINVAL, "gba_ba_eliminate without gba_ba_set_shard") : GBA_EINVAL;
  BaParams &p = c->p;
  cudaStream_t s = c->stream;
  const Status &st = *c->st_h;
  CK(cudaSetDevice(c->device));
  stage_begin(c, GBA_STAGE_FACTOR);
  // damping d += ep + lm d on the SUMMED diagonal, by linearity: own band rows in full (only this rank holds them),
  // closure rows lm d_local on every rank and ep once
  k_damp<<<(6 * p.P + 255) / 256, 256, 0, s>>>(p, p.lm, p.ep, c->seg_lo, c->seg_n, c->first_rank, st.n_seg, st.seg_tiles);
  LAUNCH_CHECK("k_damp");
  int rc = launch_chain(c, 1);
  if (rc) return rc;
  stage_end(c);
  return GBA_OK;
}

extern "C" int gba_ba_system_part(gba_ctx *c, int which, double **buf, size_t *count) {
  if (!c || !c->begun || !buf || !count) return GBA_EINVAL;
  const Status &st = *c->st_h;
  if (which == 0) {
    *buf = c->sys;
    *count = c->sys_count;
  } else {  // closure x closure tiles + the whole rhs: one contiguous range (k_tile_fill numbers those tiles last)
    *buf = c->sys + (size_t)st.cc_slot0 * GBA_NB2;
    *count = c->sys_count - (size_t)st.cc_slot0 * GBA_NB2;
  }
  return GBA_OK;
}

extern "C" int gba_ba_finish(gba_ctx *c) {
  if (!c || !c->begun || c->seg_n < 1) return c ? fail_msg(c, GBA_EINVAL, "gba_ba_finish without gba_ba_set_shard") : GBA_EINVAL;
  CK(cudaSetDevice(c->device));
  stage_begin(c, GBA_STAGE_FACTOR);
  int rc = launch_chain(c, 2);
  if (rc) return rc;
  stage_end(c);
  return solve_tail(c, true);
}

// index build + ordering + tile structure of a graph given by its edge indices alone (no per-pixel data): what a
// caller needs (GBA_GET_PERM, gba_query) to decide which rank owns which source keyframe
extern "C" int gba_plan(gba_ctx *c, const int64_t *ii, const int64_t *jj, int n_edges, int n_frames, int t0, int t1,
                        void *stream) {
  if (!c) return GBA_EINVAL;
  c->begun = 0;
  if (!ii || !jj || n_edges < 1 || n_edges > c->E_max) return fail_msg(c, GBA_EINVAL, "gba_plan: bad edge list");
  if (n_frames < 1 || n_frames > c->N_max || t0 < 0 || t1 <= t0 || t1 > n_frames)
    return fail_msg(c, GBA_EINVAL, "gba_plan: bad frame window");
  CK(cudaSetDevice(c->device));
  c->stream = (cudaStream_t)stream;
  c->launches = 0;
  BaParams &p = c->p;
  memset(&p, 0, sizeof(p));
  p.ii = ii; p.jj = jj; p.N = n_frames; p.E = n_edges; p.HW = c->HW; p.wd = c->wd; p.t0 = t0; p.t1 = t1; p.P = t1 - t0;
  p.mode = GBA_MODE_DROID_CUDA;
  p.cc_list = c->cc_list; p.kx = c->kx; p.fmap = c->fmap; p.deg = c->deg; p.eptr = c->eptr; p.eidx = c->eidx; p.ek = c->ek;
  p.goff = c->goff; p.st = c->st_d; p.tslot = c->tslot; p.perm = c->perm; p.iperm = c->iperm; p.nT = c->nT_max;
  c->last_P = p.P;
  return build_index(c, nullptr, 0, true);
}

extern "C" int gba_ba_end(gba_ctx *c, float *dx_out, float *dz_out) {
  if (!c || !c->begun) return GBA_EINVAL;
  if (dx_out)
    CK(cudaMemcpyAsync(dx_out, c->dx, (size_t)6 * c->p.P * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
  if (dz_out && !c->p.motion_only)
    CK(cudaMemcpyAsync(dz_out, c->dzbuf, (size_t)c->st_h->K * c->HW * sizeof(float), cudaMemcpyDeviceToDevice,
                       c->stream));
  return GBA_OK;
}

extern "C" int gba_ba(gba_ctx *c, float *poses, float *disps, const float *intrinsics, const float *disps_sens,
                      const float *targets, const float *weights, const float *eta, const int64_t *ii,
                      const int64_t *jj, int n_frames, int n_edges, int eta_rows, int t0, int t1, int iters, float lm,
                      float ep, int motion_only, int depth_only, int mode, float *dx_out, float *dz_out, void *stream) {
  int rc = gba_ba_begin(c, poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, n_frames, n_edges,
                        eta_rows, t0, t1, lm, ep, motion_only, depth_only, mode, nullptr, 0, nullptr, nullptr, nullptr,
                        0, stream);
  if (rc) return rc;
  for (int it = 0; it < iters; it++) {
    if ((rc = gba_ba_build(c))) return rc;
    if ((rc = gba_ba_solve(c))) return rc;
  }
  return gba_ba_end(c, dx_out, dz_out);
}

// ---- DSPO stage 2 ------------------------------------------------------------------------------
extern "C" int gba_dspo(gba_ctx *c, const float *poses, float *disps, const float *intrinsics, const float *targets,
                        const float *weights, const float *eta, const int64_t *ii, const int64_t *jj,
                        const float *mono_disps, float *scales, float *shifts, const uint8_t *valid_mask,
                        int n_frames, int n_edges, int eta_rows, int ignore_frames, float lm, float ep, float alpha,
                        int interleaved, void *stream) {
  if (!c) return GBA_EINVAL;
  c->begun = 0;
  if (!poses || !disps || !intrinsics || !targets || !weights || !eta || !ii || !jj || !mono_disps || !scales ||
      !shifts || !valid_mask)
    return fail_msg(c, GBA_EINVAL, "NULL array argument");
  if (n_frames < 1 || n_frames > c->N_max)
    return fail_msg(c, GBA_ECAPACITY, "n_frames %d outside [1,%d]", n_frames, c->N_max);
  if (n_edges < 1 || n_edges > c->E_max) return fail_msg(c, GBA_ECAPACITY, "n_edges %d outside [1,%d]", n_edges, c->E_max);
  if ((((uintptr_t)disps | (uintptr_t)targets | (uintptr_t)weights | (uintptr_t)eta | (uintptr_t)mono_disps) & 15) != 0 ||
      ((uintptr_t)valid_mask & 3) != 0)
    return fail_msg(c, GBA_EINVAL, "disps, targets, weights, eta and mono_disps must be 16-byte aligned, valid_mask 4-byte aligned");
  CK(cudaSetDevice(c->device));
  c->stream = (cudaStream_t)stream;
  c->launches = 0;
  cudaStream_t s = c->stream;
  BaParams &p = c->p;
  memset(&p, 0, sizeof(p));
  p.poses = const_cast<float *>(poses); p.disps = disps; p.intr = intrinsics; p.ii = ii; p.jj = jj;
  p.N = n_frames; p.E = n_edges; p.HW = c->HW; p.wd = c->wd; p.t0 = 0; p.t1 = n_frames; p.P = n_frames;
  p.mode = GBA_MODE_GEOM_PY;  // kx = unique(ii) (ba.py:139)
  p.cc_list = c->cc_list; p.kx = c->kx; p.fmap = c->fmap; p.deg = c->deg; p.eptr = c->eptr; p.eidx = c->eidx; p.ek = c->ek;
  p.goff = c->goff; p.st = c->st_d; p.geo = c->geo;
  int rc = build_index(c, nullptr, 0, false);
  if (rc) return rc;
  const Status &st = *c->st_h;
  if (eta_rows != st.K) return fail_msg(c, GBA_EINVAL, "eta has %d rows but unique(ii) has %d", eta_rows, st.K);
  DspoParams d;
  memset(&d, 0, sizeof(d));
  d.poses = poses; d.disps = disps; d.intr = intrinsics; d.targets = targets; d.weights = weights; d.eta = eta;
  d.mono = mono_disps; d.scales = scales; d.shifts = shifts; d.mask = valid_mask; d.ii = ii; d.jj = jj;
  d.N = n_frames; d.E = n_edges; d.HW = c->HW; d.wd = c->wd; d.ignore_frames = ignore_frames;
  d.interleaved = interleaved; d.lm = lm; d.ep = ep;
  d.sqrt_alpha = sqrtf(alpha);  // torch.tensor(alpha).sqrt() is fp32 (ba.py:141)
  d.kx = c->kx; d.eptr = c->eptr; d.eidx = c->eidx; d.st = c->st_d; d.geo = c->geo; d.Q = c->Q; d.W = c->W;
  d.part = c->dspo_part; d.dwq = c->dwq; d.bad = c->dspo_bad;
  d.nchunk = (c->HW + 511) / 512;
  stage_begin(c, GBA_STAGE_DSPO);
  CK(cudaMemsetAsync(c->dspo_bad, 0, sizeof(int), s));
  k_edge_geo<<<(p.E + 255) / 256, 256, 0, s>>>(p);
  LAUNCH_CHECK("k_edge_geo");
  dim3 grid(d.nchunk, st.K);
  if (interleaved) k_dspo_lin<128, true><<<grid, 128, 0, s>>>(d);
  else k_dspo_lin<128, false><<<grid, 128, 0, s>>>(d);
  LAUNCH_CHECK("k_dspo_lin");
  k_dspo_solve<<<1, 256, 0, s>>>(d);
  LAUNCH_CHECK("k_dspo_solve");
  k_dspo_update<128><<<grid, 128, 0, s>>>(d);
  LAUNCH_CHECK("k_dspo_update");
  if (st.K < n_frames) {
    k_clamp_rest<<<dim3((c->HW / 4 + 127) / 128, n_frames), 128, 0, s>>>(disps, c->fmap, c->HW);
    LAUNCH_CHECK("k_clamp_rest");
  }
  stage_end(c);
  return GBA_OK;
}

// ---- stage timing -------------------------------------------------------------------------------
extern "C" int gba_set_timing(gba_ctx *c, int enable) {
  if (!c) return GBA_EINVAL;
  c->timing = enable ? 1 : 0;
  c->ev_used = 0;
  c->ev_stage->clear();
  c->ev_launch->clear();
  return GBA_OK;
}

extern "C" int gba_get_timing(gba_ctx *c, float *ms, int *launches) {
  if (!c || !ms) return GBA_EINVAL;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < GBA_N_STAGES; i++) {
    ms[i] = 0.f;
    if (launches) launches[i] = 0;
  }
  for (size_t i = 0; i < c->ev_stage->size(); i++) {
    float t = 0.f;
    CK(cudaEventElapsedTime(&t, (*c->ev_pool)[2 * i], (*c->ev_pool)[2 * i + 1]));
    ms[(*c->ev_stage)[i]] += t;
    if (launches) launches[(*c->ev_stage)[i]] += (*c->ev_launch)[i];
  }
  c->ev_used = 0;
  c->ev_stage->clear();
  c->ev_launch->clear();
  return GBA_OK;
}

// ---- introspection ------------------------------------------------------------------------------
extern "C" int gba_last_depth_frames(const gba_ctx *c) { return (c && c->st_h) ? c->st_h->K : GBA_EINVAL; }

extern "C" int gba_query(gba_ctx *c, gba_info *info) {
  if (!c || !info) return GBA_EINVAL;
  CK(cudaSetDevice(c->device));
  CK(cudaMemcpyAsync(c->st_h, c->st_d, sizeof(Status), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  const Status &st = *c->st_h;
  info->n_depth_frames = st.K;
  info->n_poses = c->last_P;
  info->max_out_degree = st.max_deg;
  info->chol_failures = st.chol_fail;
  info->n_tiles = st.n_tiles;
  info->n_active_tiles = st.n_slots;
  info->kernel_launches = c->launches;
  // sequential tile columns of the factorisation: one band segment's worth plus the closure block
  info->chain_columns = (st.n_seg > 1 ? st.seg_tiles : 0) + st.seg2_tiles +
                        (st.n_tiles - (st.n_seg > 1 ? st.n_seg * st.seg_tiles : 0) - st.n_seg2 * st.seg2_tiles);
  info->last_fail_code = st.last_fail;
  info->n_segments = st.n_seg;
  info->segment_tiles = st.seg_tiles;
  int wd0 = 0;
  CK(cudaMemcpy(&wd0, c->wdog, sizeof(int), cudaMemcpyDeviceToHost));
  info->watchdog_trips = wd0;
  return GBA_OK;
}

extern "C" int gba_get(gba_ctx *c, int what, void *dst, size_t bytes) {
  if (!c || !dst) return GBA_EINVAL;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  const Status &st = *c->st_h;
  const BaParams &p = c->p;
  const void *src = nullptr;
  size_t need = 0;
  switch (what) {
    case GBA_GET_KX: src = c->kx; need = (size_t)st.K * 4; break;
    case GBA_GET_EPTR: src = c->eptr; need = (size_t)(st.K + 1) * 4; break;
    case GBA_GET_EIDX: src = c->eidx; need = (size_t)p.E * 4; break;
    case GBA_GET_EDGE_HV: src = c->edge_hv; need = (size_t)p.E * 32 * 4; break;
    case GBA_GET_Q: src = c->Q; need = (size_t)st.K * p.HW * 4; break;
    case GBA_GET_W: src = c->W; need = (size_t)st.K * p.HW * 4; break;
    case GBA_GET_GRAM: src = c->gram; need = (size_t)st.gram_floats * 4; break;
    case GBA_GET_WATCHDOG: src = c->wdog; need = 256 * 4; break;
    case GBA_GET_PERM: src = c->perm; need = (size_t)c->last_P * 4; break;
    case GBA_GET_CHAIN_TRACE:
      if (!c->sb.trace) return fail_msg(c, GBA_EINVAL, "gba_get: no trace buffer (create the context with GBA_CHAIN_TRACE=1 set)");
      src = c->sb.trace; need = std::min(bytes, (1 + 4 * (size_t)CHAIN_TRACE_CAP) * sizeof(unsigned long long)); break;
    case GBA_GET_TILEMASK: {
      const int nTa = st.n_tiles;
      need = (size_t)nTa * nTa;
      if (bytes < need) return fail_msg(c, GBA_EINVAL, "gba_get: buffer too small (%zu < %zu)", bytes, need);
      std::vector<unsigned int> bits((size_t)nTa * TS_WORDS);
      CK(cudaMemcpy(bits.data(), c->sb.tbits, bits.size() * sizeof(unsigned int), cudaMemcpyDeviceToHost));
      unsigned char *out = (unsigned char *)dst;
      for (int i = 0; i < nTa; i++)
        for (int j = 0; j < nTa; j++) out[(size_t)i * nTa + j] = (bits[(size_t)i * TS_WORDS + (j >> 5)] >> (j & 31)) & 1;
      return GBA_OK;
    }
    case GBA_GET_SYSTEM: {
      const size_t n = 6 * (size_t)p.P;
      need = (n * n + n) * sizeof(double);
      if (bytes < need) return fail_msg(c, GBA_EINVAL, "gba_get: buffer too small (%zu < %zu)", bytes, need);
      std::vector<double> h(c->sys_count);
      std::vector<int> slot((size_t)p.nT * p.nT), perm(p.P);
      CK(cudaMemcpy(h.data(), c->sys, c->sys_count * sizeof(double), cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(slot.data(), c->tslot, slot.size() * sizeof(int), cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(perm.data(), c->perm, perm.size() * sizeof(int), cudaMemcpyDeviceToHost));
      double *out = (double *)dst;
      memset(out, 0, need);
      // undo the arrowhead ordering: element (r, q) of the caller's ordering lives at the permuted
      // position, on whichever side of the diagonal is stored
      for (size_t r = 0; r < n; r++)
        for (size_t q = 0; q <= r; q++) {
          size_t pr = 6 * (size_t)perm[r / 6] + r % 6, pq = 6 * (size_t)perm[q / 6] + q % 6;
          if (pr < pq) std::swap(pr, pq);
          const int sl = slot[(pr / GBA_NB) * p.nT + pq / GBA_NB];
          if (sl >= 0) out[r * n + q] = h[(size_t)sl * GBA_NB2 + (pr % GBA_NB) * GBA_NB + pq % GBA_NB];
        }
      for (size_t r = 0; r < n; r++) out[n * n + r] = h[(size_t)st.n_slots * GBA_NB2 + 6 * (size_t)perm[r / 6] + r % 6];
      return GBA_OK;
    }
    default: return fail_msg(c, GBA_EINVAL, "gba_get: unknown selector %d", what);
  }
  if (bytes < need) return fail_msg(c, GBA_EINVAL, "gba_get: buffer too small (%zu < %zu)", bytes, need);
  CK(cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
  return GBA_OK;
}
